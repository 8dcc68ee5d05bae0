"""CPU-side checks: the C ABI exports what the header declares, host logic (bucket assignment, reducer
state machine, stores) matches the oracle / the reference's literal vectors, and compute entry points
fail loudly without a GPU instead of falling back."""
import ctypes as C
import json
import os
import re
import threading

import numpy as np
import pytest

from oracle import oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    return prb


def test_library_exports_every_declared_symbol(prb):
    header = open(os.path.join(ROOT, "include", "paddle_ray_b200.h")).read()
    declared = set(re.findall(r"\b(prb_[a-z0-9_]+)\s*\(", header))
    assert len(declared) > 50
    from paddle_ray_b200 import _lib
    lib = C.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), "%s is declared in the header but not exported" % name
    # and the ctypes binding covers all of them
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


def test_header_cites_the_reference_for_every_operator():
    header = open(os.path.join(ROOT, "include", "paddle_ray_b200.h")).read()
    for op in ("prb_adamw", "prb_adam", "prb_merged_adam", "prb_fused_adam", "prb_check_finite_and_unscale",
               "prb_squared_l2_norm", "prb_scale", "prb_concat", "prb_assign_group_by_size", "prb_reducer_create",
               "prb_comm_all_reduce", "prb_update_loss_scaling"):
        idx = header.index("int32_t %s(" % op)
        assert "ref:" in header[max(0, idx - 1500):idx], op


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="GPU present")
def test_no_cpu_fallback(prb):
    from paddle_ray_b200._lib import PRBError
    assert prb.device_count() == 0
    host = prb.Tensor(16, "float32", device=-1)
    one = prb.Tensor(1, "float32", device=-1)
    with pytest.raises(PRBError, match="no CPU fallback"):
        prb._C_ops.adamw_(host, host, one, host, host, one, one, None, None, 0.9, 0.999, 1e-8, 1.0, 0.01, True)
    with pytest.raises(PRBError, match="no CPU fallback"):
        prb._C_ops.scale_(host, 0.5)
    with pytest.raises(PRBError, match="no CPU fallback"):
        prb.Tensor(16, "float32", device=0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "paddle_ray_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("oracle/dp_tail_oracle.c", "").replace("the CPU oracle", "") \
                    .replace("(oracle/", "("), f


class _T:
    def __init__(self, numel, dtype):
        self.numel, self.dtype = numel, dtype


def test_assign_group_by_size_golden(prb, golden_dir):
    names = {"float32": orc.F32, "float64": orc.F64, "float16": orc.F16, "bfloat16": orc.BF16}
    cases = json.load(open(os.path.join(golden_dir, "bucket_cases.json")))
    for c in cases:
        ts = [_T(int(np.prod(s)), names[d]) for s, d in zip(c["shapes"], c["dtypes"])]
        got = prb.distributed.assign_group_by_size(ts, c["is_sparse"], c["limits"], c["tensor_indices"])
        assert got == c["expected"], c["name"]


def test_assign_group_by_size_random_vs_oracle(prb):
    rng = np.random.default_rng(0)
    for trial in range(200):
        n = int(rng.integers(1, 40))
        numels = [int(x) for x in rng.integers(1, 5000, n)]
        dts = [int(rng.choice([orc.F32, orc.BF16, orc.F16, orc.F64])) for _ in range(n)]
        sparse = [bool(rng.random() < 0.1) for _ in range(n)]
        limits = [int(x) for x in rng.integers(0, 20000, int(rng.integers(1, 4)))]
        ti = [int(x) for x in rng.permutation(n)] if trial % 3 == 0 else []
        want = orc.assign_group_by_size(numels, dts, sparse, limits, ti)
        got = prb.distributed.assign_group_by_size([_T(a, b) for a, b in zip(numels, dts)], sparse, limits, ti)
        assert got == want


def test_assign_group_by_size_model_configs(prb):
    """ResNet-50 fp32 -> 1 MiB first bucket then 25 MiB buckets; a bucket may overshoot by one tensor."""
    from bench_configs import param_shapes
    shapes = param_shapes("resnet50")
    ts = [_T(int(np.prod(s)), orc.F32) for s in shapes]
    groups = prb.distributed.assign_group_by_size(ts, [False] * len(ts), [1 << 20, 25 << 20])
    assert sorted(v for g in groups for v in g) == list(range(len(ts)))
    sizes = [sum(ts[v].numel * 4 for v in g) for g in groups]
    assert sizes[0] >= (1 << 20) and all(s >= (25 << 20) for s in sizes[1:-1])
    assert groups == orc.assign_group_by_size([t.numel for t in ts], [orc.F32] * len(ts), [False] * len(ts),
                                              [1 << 20, 25 << 20])


def test_assign_group_by_size_errors(prb):
    from paddle_ray_b200._lib import PRBError
    with pytest.raises(PRBError, match="permutation"):
        prb.distributed.assign_group_by_size([_T(1, orc.F32), _T(1, orc.F32)], [False, False], [400], [0, 0])
    with pytest.raises(PRBError, match="is_sparse_gradient"):
        prb.distributed.assign_group_by_size([_T(1, orc.F32)], [False, False], [400])


def _host_grads(prb, sizes, dtype="float32"):
    return [prb.Tensor(n, dtype, device=-1) for n in sizes]


def test_reducer_state_machine_matches_model_on_random_orders(prb):
    rng = np.random.default_rng(1)
    for trial in range(50):
        n = int(rng.integers(2, 30))
        sizes = [int(x) for x in rng.integers(1, 3000, n)]
        ts = [_T(s, orc.F32) for s in sizes]
        groups = prb.distributed.assign_group_by_size(ts, [False] * n, [2000, 6000])
        groups = list(reversed(groups))
        red = prb.distributed.EagerReducer(_host_grads(prb, sizes), groups, [False] * n, None)
        model = orc.ReducerModel(groups, n)
        for step in range(2):
            red.prepare_for_backward()
            model.prepare_for_backward()
            order = list(rng.permutation(n)) if trial % 2 else list(reversed(range(n)))
            for v in order:
                red.mark_var_ready(int(v))
                model.mark_var_ready(int(v))
                assert red.launched_groups() == model.launched
                assert red.finalized == model.finalized
            assert red.launched_groups() == list(range(len(groups))) and red.finalized


def test_reducer_error_behaviour(prb):
    from paddle_ray_b200._lib import PRBError
    sizes = [10, 20, 30, 40]
    groups = [[3, 2], [1, 0]]
    red = prb.distributed.EagerReducer(_host_grads(prb, sizes), groups, [False] * 4, None)
    red.prepare_for_backward()
    red.mark_var_ready(3)
    with pytest.raises(PRBError, match="has been ready before"):   # reducer.cc:779-811
        red.mark_var_ready(3)
    with pytest.raises(PRBError, match="Out of bounds variable index"):  # reducer.cc:731-737
        red.mark_var_ready(99)
    with pytest.raises(PRBError, match="A serious error has occurred here"):  # reducer.cc:683-698
        red.prepare_for_backward()
    for v in (2, 1, 0):
        red.mark_var_ready(v)
    assert red.finalized
    red.mark_var_ready(0)  # after finalize hooks only re-alias: a no-op (reducer.cc:740-758)
    red.prepare_for_backward()
    # dtype uniformity inside a group (reducer.cc:607-612), empty groups (:536-540), zero-size tensors (:589-593)
    mixed = [prb.Tensor(4, "float32", device=-1), prb.Tensor(4, "bfloat16", device=-1)]
    with pytest.raises(PRBError, match="unexpected dtype"):
        prb.distributed.EagerReducer(mixed, [[0, 1]], [False, False], None)
    with pytest.raises(PRBError, match="elements is 0"):
        prb.distributed.EagerReducer(_host_grads(prb, [0, 4]), [[0, 1]], [False, False], None)
    with pytest.raises(PRBError, match="find_unused_parameters=False"):
        red.prepare_for_backward([1])


def test_reducer_unused_vars_are_marked_with_the_first_hook(prb):
    sizes = [10, 20, 30, 40]
    groups = [[3, 2], [1, 0]]
    red = prb.distributed.EagerReducer(_host_grads(prb, sizes), groups, [False] * 4, None, None, True)
    red.prepare_for_backward(unused_vars=[2, 0])
    red.mark_var_ready(3)          # marks 2 and 0 ready too -> group 0 complete
    assert red.launched_groups() == [0]
    red.mark_var_ready(1)
    assert red.launched_groups() == [0, 1] and red.finalized


def test_stores_roundtrip(tmp_path):
    from paddle_ray_b200.distributed import FileStore, TCPStore
    fs = FileStore(str(tmp_path / "fs"), timeout=5)
    fs.set("ProcessGroupNCCL/nccl_ids/0/0", b"\x01" * 128)
    assert fs.get("ProcessGroupNCCL/nccl_ids/0/0") == b"\x01" * 128
    assert fs.add("ctr", 2) == 2 and fs.add("ctr", 3) == 5
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    master = TCPStore("127.0.0.1", port, True, 2, timeout=10)
    got = {}

    def peer():
        c = TCPStore("127.0.0.1", port, False, 2, timeout=10)
        got["v"] = c.get("k")       # blocks until the master publishes
        got["n"] = c.add("n", 1)
        c.close()

    t = threading.Thread(target=peer)
    t.start()
    master.set("k", b"unique-id-bytes")
    t.join(10)
    assert got["v"] == b"unique-id-bytes" and got["n"] == 1
    assert master.add("n", 1) == 2
    master.close()


def test_parallel_env_reads_reference_and_torchrun_names(monkeypatch):
    from paddle_ray_b200.distributed import ParallelEnv
    for k in ("PADDLE_TRAINER_ID", "PADDLE_TRAINERS_NUM", "PADDLE_MASTER", "FLAGS_selected_gpus", "RANK", "WORLD_SIZE",
              "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("PADDLE_TRAINER_ID", "3")
    monkeypatch.setenv("PADDLE_TRAINERS_NUM", "8")
    monkeypatch.setenv("PADDLE_MASTER", "10.0.0.1:6170")
    monkeypatch.setenv("FLAGS_selected_gpus", "5")
    e = ParallelEnv()
    assert (e.rank, e.world_size, e.device_id, e.master_addr, e.master_port) == (3, 8, 5, "10.0.0.1", 6170)
    for k in ("PADDLE_TRAINER_ID", "PADDLE_TRAINERS_NUM", "PADDLE_MASTER", "FLAGS_selected_gpus"):
        monkeypatch.delenv(k)
    monkeypatch.setenv("RANK", "1")
    monkeypatch.setenv("WORLD_SIZE", "2")
    monkeypatch.setenv("LOCAL_RANK", "1")
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", "29501")
    e = ParallelEnv()
    assert (e.rank, e.world_size, e.device_id, e.master_port, e.under_torchrun) == (1, 2, 1, 29501, True)


def test_fleet_build_groups_matches_reference_rule(prb):
    """parallel.py:136-153: the size check comes BEFORE adding, and a dtype change always starts a new group."""
    mk = lambda n, dt: prb.Tensor(n, dt, device=-1)  # noqa: E731
    vs = [mk(100, "float32"), mk(100, "float32"), mk(100, "float32"), mk(50, "bfloat16"), mk(50, "bfloat16"), mk(10, "float32")]
    groups = prb.distributed.build_groups(vs, 700)
    assert [[v.numel for v in g] for g in groups] == [[100, 100], [100], [50, 50], [10]]
    assert [[v.numel for v in g] for g in prb.distributed.build_groups(vs, 10 ** 9)] == [[100, 100, 100], [50, 50], [10]]


def test_sharding_optimizer_partitions_parameters_greedily():
    """ref dygraph_sharding_optimizer.py:101-126: each parameter goes to the rank holding the fewest elements so far."""
    from paddle_ray_b200.core import Tensor
    from paddle_ray_b200.distributed.sharding import DygraphShardingOptimizer

    class HCG:
        def get_sharding_parallel_world_size(self):
            return 3

        def get_sharding_parallel_rank(self):
            return 1

        def get_sharding_parallel_group(self):
            return None

    class Inner:
        def __init__(self, parameters, **kw):
            self.parameters, self.kw, self._grad_clip = parameters, kw, None

    sizes = [100, 10, 50, 60, 5, 200, 1]
    params = [Tensor(n, "float32", device=-1, name="p%d" % i) for i, n in enumerate(sizes)]
    opt = DygraphShardingOptimizer(HCG(), None, params, Inner, learning_rate=0.1)
    got = {r: [p.name for p in ps] for r, ps in opt._rank2params.items()}
    # p0->0 (100,0,0) p1->1 (100,10,0) p2->2 (100,10,50) p3->1 (100,70,50) p4->2 (100,70,55) p5->2 (100,70,255) p6->1
    assert got == {0: ["p0"], 1: ["p1", "p3", "p6"], 2: ["p2", "p4", "p5"]}
    assert [p.name for p in opt._inner_optimizer.parameters] == ["p1", "p3", "p6"] and opt.kw == {"learning_rate": 0.1}
    assert opt._param2rank["p5"] == 2
    with pytest.raises(TypeError):
        DygraphShardingOptimizer(HCG(), None, tuple(params), Inner)


def test_reducer_sparse_groups_state_machine(prb):
    """A sparse variable is a group of its own (reducer.cc:94-99, 543-551); it needs a SelectedRows gradient when its
    hook fires (:858-888); the backward is finalized once the dense groups are launched AND the sparse groups are
    exchanged."""
    from paddle_ray_b200._lib import PRBError
    sizes = [10, 20, 30, 40]
    ts = [_T(s, orc.F32) for s in sizes]
    is_sparse = [False, True, False, False]
    groups = prb.distributed.assign_group_by_size(ts, is_sparse, [400, 400])
    assert [1] in groups
    grads = _host_grads(prb, sizes)
    grads[1] = None
    red = prb.distributed.EagerReducer(grads, list(reversed(groups)), is_sparse, None)
    assert red.dense_group_count == len(groups) - 1
    holder = {"g": None}
    red.set_sparse_source(1, lambda: holder["g"])
    red.prepare_for_backward()
    with pytest.raises(PRBError, match="should have gradient"):
        red.mark_var_ready(1)
    holder["g"] = grads[0]
    with pytest.raises(PRBError, match="must have a selectedrows gradient"):
        red.mark_var_ready(1)
    holder["g"] = prb.SelectedRows([2, 0, 2], 5, prb.Tensor(3 * 4, "float32", device=-1))
    assert holder["g"].width == 4
    for v in (3, 2):
        red.mark_var_ready(v)
    assert not red.finalized
    red.mark_var_ready(1)
    assert not red.finalized          # a dense variable is still missing
    with pytest.raises(PRBError, match="marked ready twice"):
        red.mark_var_ready(1)
    red.mark_var_ready(0)
    assert red.finalized and red.launched_groups() == list(range(red.dense_group_count))
    red.prepare_for_backward()
    assert not red.finalized
    with pytest.raises(PRBError, match="own"):
        prb.distributed.EagerReducer(grads, [[0, 1], [2, 3]], is_sparse, None)
    # sparse variables only: no native reducer at all
    only = prb.distributed.EagerReducer([None], [[0]], [True], None)
    only.set_sparse_source(0, lambda: holder["g"])
    only.prepare_for_backward()
    only.mark_vars_ready([0])
    assert only.finalized


def test_shard_ranges_cover_every_element_once(prb):
    """prb_reducer_shard_ranges on the host-only reducer (one rank owns everything): ranges follow the arena order
    (groups in launch order, slots in order), one range per variable, all elements covered."""
    sizes = [10, 20000, 8192, 1, 30000]
    groups = [[4, 3], [2], [1, 0]]
    red = prb.distributed.EagerReducer(_host_grads(prb, sizes), groups, [False] * 5, None)
    got = red.shard_ranges(0)
    assert got == [(4, 0, 30000), (3, 0, 1), (2, 0, 8192), (1, 0, 20000), (0, 0, 10)]
    from paddle_ray_b200._lib import PRBError
    with pytest.raises(PRBError):
        red.shard_ranges(1)
    # any world size, both layouts: every element owned exactly once, one range per variable and rank, chunk-aligned
    # offsets; equal shares to within a chunk -- of the whole arena, or of every bucket
    rng = np.random.default_rng(4)
    for trial in range(20):
        n = int(rng.integers(3, 40))
        sizes = [int(x) for x in rng.integers(1, 60000, n)]
        groups = list(reversed(prb.distributed.assign_group_by_size([_T(s, orc.F32) for s in sizes], [False] * n, [60000, 200000])))
        for per_bucket in (False, True):
            red = prb.distributed.EagerReducer(_host_grads(prb, sizes), groups, [False] * n, None)
            red.set_shard_layout(per_bucket)
            for world in (2, 3, 8):
                owned = np.zeros(sum(sizes), np.int32)
                start = np.concatenate([[0], np.cumsum(sizes)])
                per_rank = []
                for r in range(world):
                    rr = red.shard_ranges(r, world)
                    assert len({v for v, _, _ in rr}) == len(rr)
                    for v, off, ln in rr:
                        assert off % 8192 == 0 and ln > 0 and off + ln <= sizes[v]
                        owned[start[v] + off:start[v] + off + ln] += 1
                    per_rank.append(rr)
                assert (owned == 1).all()
                if per_bucket:
                    for g in groups:
                        ch = [sum((ln + 8191) // 8192 for v, _, ln in rr if v in g) for rr in per_rank]
                        assert max(ch) - min(ch) <= 1, (g, ch)
                else:
                    ch = [sum((ln + 8191) // 8192 for _, _, ln in rr) for rr in per_rank]
                    assert max(ch) - min(ch) <= 1, ch


def test_gather_kernel_chunk_walk_restated():
    """The index arithmetic of nvl_gather_chunks_kernel (csrc/prb_nvlink.cu: CTA b serves peer (rank + 1 + b mod (n-1))
    mod n and walks that peer's chunks of every ownership segment as one list), restated in Python: for every rank,
    grid size and segment table, the CTAs that serve peer q visit exactly q's chunks -- the n-th of every segment
    that prb_reducer_shard_ranges assigns to q -- once each, and never the rank's own."""
    import random

    def walk(nseg, seg_first, n, r, grid):
        peers, got = n - 1, {}
        for b in range(grid):
            pi = b % peers
            q = (r + 1 + pi) % n
            fi, step = b // peers, (grid - pi + peers - 1) // peers
            seg = seg_base = 0
            s0 = seg_first[0] + (seg_first[1] - seg_first[0]) * q // n
            s1 = seg_first[0] + (seg_first[1] - seg_first[0]) * (q + 1) // n
            while True:
                while seg < nseg and fi - seg_base >= s1 - s0:
                    seg_base += s1 - s0
                    seg += 1
                    if seg < nseg:
                        f, ln = seg_first[seg], seg_first[seg + 1] - seg_first[seg]
                        s0, s1 = f + ln * q // n, f + ln * (q + 1) // n
                if seg >= nseg:
                    break
                got.setdefault(q, []).append(s0 + (fi - seg_base))
                fi += step
        return got

    random.seed(1)
    for trial in range(200):
        n = random.choice([2, 3, 4, 8])
        nseg = random.randint(1, 30)
        lens = [random.randint(0, 3 if trial % 3 == 0 else 40) for _ in range(nseg)]
        sf = [0]
        for ln in lens:
            sf.append(sf[-1] + ln)
        for r in range(n):
            peers = n - 1
            ctas = (sf[-1] + n - 1) // n * peers                    # nvl_gather_params' grid
            cap = random.choice([296, 64, 14])
            ctas = max(peers, cap // peers * peers if ctas > cap else ctas)
            got = walk(nseg, sf, n, r, ctas)
            assert r not in got
            for q in range(n):
                if q != r:
                    want = [c for s in range(nseg)
                            for c in range(sf[s] + (sf[s + 1] - sf[s]) * q // n, sf[s] + (sf[s + 1] - sf[s]) * (q + 1) // n)]
                    assert sorted(got.get(q, [])) == want
