"""World-size-2 CPU tests (gloo): the N>1 host logic that needs no GPU.

 * the NCCL unique-id hand-off protocol of init_parallel_env (rank 0 publishes under the reference's key, peers
   fetch) over the package's TCPStore, between two real processes;
 * the oracle's data-parallel gradient path (concat -> scale(1/n) -> allreduce -> split) with a gloo allreduce:
   grads equal across ranks and equal to the scaled sum -- the CPU twin of tests/dist_worker.py, mirroring the
   reference's test_parallel_dygraph_dataparallel_cpuonly.py.
"""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, store_port, gloo_port, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import torch

    from oracle import oracle as orc
    from paddle_ray_b200.distributed import TCPStore

    # (1) unique-id hand-off
    store = TCPStore("127.0.0.1", store_port, rank == 0, world, timeout=60)
    key = "ProcessGroupNCCL/nccl_ids/0/0"
    if rank == 0:
        uid = bytes(range(128))
        store.set(key, uid)
    else:
        uid = store.get(key)
    assert uid == bytes(range(128))
    arrived = store.add("arrived", 1)
    assert 1 <= arrived <= world

    # (2) oracle DP path with gloo allreduce
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % gloo_port, rank=rank, world_size=world)
    sizes = [64, 1000, 7, 4097]
    grads = [np.random.default_rng(1000 * rank + i).normal(0, 1e-2, n).astype(np.float32) for i, n in enumerate(sizes)]
    bucket = orc.scale(orc.concat(grads), orc.F32, 1.0 / world)
    t = torch.from_numpy(bucket)
    dist.all_reduce(t)  # in place, sum
    outs = orc.split(bucket, sizes)
    for i, o in enumerate(outs):
        np.save(os.path.join(out_dir, "g%d_r%d.npy" % (i, rank)), o)
        want = sum(orc.scale(np.random.default_rng(1000 * r + i).normal(0, 1e-2, sizes[i]).astype(np.float32), orc.F32,
                             1.0 / world).astype(np.float64) for r in range(world))
        np.testing.assert_array_equal(o, want.astype(np.float32))  # two addends: order cannot matter
    # max-over-ranks timing reduction used by bench.py
    tm = torch.tensor([float(rank + 1)])
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    assert tm.item() == float(world)
    dist.barrier()
    dist.destroy_process_group()
    store.close()


def test_world_size_2_gloo(tmp_path):
    import torch.multiprocessing as mp
    store_port, gloo_port = _free_port(), _free_port()
    ctx = mp.get_context("spawn")
    procs = [ctx.Process(target=_worker, args=(r, 2, store_port, gloo_port, str(tmp_path))) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    for i in range(4):
        a = np.load(os.path.join(str(tmp_path), "g%d_r0.npy" % i))
        b = np.load(os.path.join(str(tmp_path), "g%d_r1.npy" % i))
        assert a.tobytes() == b.tobytes()


def test_bench_reference_arm_runs_gloo_allreduce_with_all_cores():
    """`bench.py --impl reference --gpus 2`: two local CPU processes, gloo allreduce per bucket, an explicit thread
    count (not the OMP_NUM_THREADS=1 a launcher exports), and the `config` keys of the GPU arm."""
    import json
    import subprocess

    # launched exactly like the driver launches it: under torch.distributed.run, whose environment
    # (TORCHELASTIC_USE_AGENT_STORE, OMP_NUM_THREADS=1, ...) must not leak into the arm's own gloo workers
    port = _free_port()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"),
                          "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1", "--workload", "resnet18"],
                         capture_output=True, timeout=240)
    assert out.returncode == 0, out.stderr.decode()[-2000:]
    env = {**os.environ, "OMP_NUM_THREADS": "1", "RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"}
    lines = [ln for ln in out.stdout.decode().splitlines() if ln.startswith("{")]
    assert len(lines) == 1, out.stdout.decode()[-1000:]   # rank 0 alone prints, ONE line
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert "gloo" in line["details"]["allreduce"]
    assert line["cpu_baseline"]["cores"] >= 2 and line["cpu_baseline"]["kind"] == "port"
    sys.path.insert(0, ROOT)
    import bench
    want = bench.config_block("x", 1, 1, 1, "float32", "float32", False, False, 36, 2)
    assert set(line["config"]) == set(want)
    # a rank other than 0 under the launcher prints nothing and exits 0
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         env={**env, "RANK": "1"}, capture_output=True, timeout=60)
    assert out.returncode == 0 and out.stdout.decode().strip() == ""
