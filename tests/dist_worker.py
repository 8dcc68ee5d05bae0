"""One rank of the multi-GPU parity check (launched by tests/test_gpu_multi.py or by hand under torchrun).

Mirrors the reference's parallel_dygraph_gradient_check_in_eager_mode.py:139-150: gradients must be EQUAL
across ranks after the sync; here additionally equal to the oracle's scale-then-sum for 2 ranks (order cannot
matter with two addends) and within tolerance for more.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import paddle_ray_b200 as prb  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def main(out_dir):
    group = prb.distributed.init_parallel_env()
    rank, n = group.rank, group.nranks
    sizes = [64, 64, 3 * 3 * 64 * 16, 1000, 7, 4096, 4097, 250000, 1, 300000]
    for dtype in (orc.F32, orc.BF16):
        params, grads_all = [], []
        for i, sz in enumerate(sizes):
            p = prb.Tensor.from_numpy(orc.from_f32(np.random.default_rng(100 + i).normal(0, 0.02, sz).astype(np.float32), dtype),
                                      dtype, name="w_%d" % i)
            per_rank = [orc.from_f32(np.random.default_rng(1000 * r + i).normal(0, 1e-2, sz).astype(np.float32), dtype)
                        for r in range(n)]
            p.grad = prb.Tensor.from_numpy(per_rank[rank], dtype)
            params.append(p)
            grads_all.append(per_rank)
        dp = prb.DataParallel(params, comm_buffer_size=0.4, last_comm_buffer_size=0.01)
        assert len(dp.group_indices) >= 3, dp.group_indices
        assert dp._reducer.allreduce_backend == os.environ.get("PRB_ALLREDUCE", "nvlink"), dp._reducer.allreduce_backend
        opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, multi_precision=dtype != orc.F32)
        for step in range(2):
            dp.prepare_for_backward()
            # hooks fire one by one, in reverse order, like a backward pass (not the batch call)
            for i in reversed(range(len(params))):
                dp.mark_var_ready(i)
            assert dp.finalized
            prb.current_stream().synchronize()
            for i, p in enumerate(params):
                got = p.grad.numpy()
                scaled = [orc.to_f32(orc.scale(g, dtype, 1.0 / n), dtype).astype(np.float64) for g in grads_all[i]]
                want = np.sum(scaled, axis=0)
                if dtype == orc.F32 and n == 2:
                    np.testing.assert_array_equal(got, want.astype(np.float32))
                else:
                    tol = 1e-6 if dtype == orc.F32 else 1e-2
                    np.testing.assert_allclose(orc.to_f32(got, dtype), want, rtol=tol, atol=tol * 1e-2)
                np.save(os.path.join(out_dir, "grad_%d_%d_%d_r%d.npy" % (dtype, step, i, rank)), got)
            opt.step()  # consumes the synchronised gradients
            prb.current_stream().synchronize()
            for i, p in enumerate(params):  # next step: every rank feeds its own fresh gradient again
                p.grad.copy_from_numpy(grads_all[i][rank])
        prb.current_stream().synchronize()
        for i, p in enumerate(params):
            np.save(os.path.join(out_dir, "param_%d_%d_r%d.npy" % (dtype, i, rank)), p.numpy())
    # pipelined consumption: the optimizer reads the bucket arena and waits per bucket (no clip, no AMP)
    params, grads_all, host = [], [], []
    for i, sz in enumerate(sizes):
        p0 = np.random.default_rng(100 + i).normal(0, 0.02, sz).astype(np.float32)
        p = prb.Tensor.from_numpy(p0, name="w_%d" % i)
        per_rank = [np.random.default_rng(1000 * r + i).normal(0, 1e-2, sz).astype(np.float32) for r in range(n)]
        p.grad = prb.Tensor.from_numpy(per_rank[rank])
        params.append(p)
        grads_all.append(per_rank)
        host.append({"p": p0.copy(), "m1": np.zeros(sz, np.float32), "m2": np.zeros(sz, np.float32)})
    dp = prb.DataParallel(params, comm_buffer_size=0.4, last_comm_buffer_size=0.01, consume_from_buckets=True)
    assert dp.pipelined
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, weight_decay=0.01)
    b1p, b2p = np.float32(0.9), np.float32(0.999)
    for step in range(3):
        dp.prepare_for_backward()
        dp.mark_all_ready()
        opt.step()
        prb.current_stream().synchronize()
        for i, (p, h) in enumerate(zip(params, host)):
            avg = np.sum([orc.scale(g, orc.F32, 1.0 / n).astype(np.float64) for g in grads_all[i]], axis=0).astype(np.float32)
            if n == 2:  # two addends: the allreduce result is exactly this
                orc.adam(h["p"], avg, h["m1"], h["m2"], None, dtype_param=orc.F32, dtype_grad=orc.F32, lr=1e-3, beta1=0.9,
                         beta2=0.999, epsilon=1e-8, coeff=0.01, use_decay=True, beta1_pow=b1p, beta2_pow=b2p)
                np.testing.assert_array_equal(p.numpy(), h["p"])
            np.save(os.path.join(out_dir, "pipe_param_%d_%d_r%d.npy" % (step, i, rank)), p.numpy())
        b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
    assert opt._beta1_pow.numpy()[0] == b1p
    assert not dp._reducer.comm_error()
    # fleet-style non-overlapped sync (hybrid_parallel_util.py:73-111): coalesce -> scale -> allreduce -> split
    fl = []
    for i, sz in enumerate(sizes[:6]):
        p = prb.Tensor.zeros(sz, "float32", name="f_%d" % i)
        p.grad = prb.Tensor.from_numpy(grads_all[i][rank])
        fl.append(p)
    prb.distributed.fused_allreduce_gradients_with_group(fl, group, bucket_size=1 << 14)
    prb.current_stream().synchronize()
    for i, p in enumerate(fl):
        want = np.sum([orc.scale(g, orc.F32, 1.0 / n, 0.0, True).astype(np.float64) for g in grads_all[i]], axis=0)
        if n == 2:
            np.testing.assert_array_equal(p.grad.numpy(), want.astype(np.float32))
        else:
            np.testing.assert_allclose(p.grad.numpy(), want, rtol=1e-6, atol=1e-8)
        np.save(os.path.join(out_dir, "fleet_grad_%d_r%d.npy" % (i, rank)), p.grad.numpy())
    # AMP bf16 + fp32 master + global-norm clip across ranks, an Inf on ONE rank only (every rank must skip, the loss
    # scale must move identically), batch and per-bucket exchange, gradients packed and produced in the arena;
    # state bit-equal across ranks, N = 2 bit-exact vs the oracle  (the same routine bench.py runs before timing)
    import bench
    for arena in (False, True):
        verdict = bench.parity_preamble(prb, group, arena=arena, save_dir=out_dir)
        assert verdict == "ok", "AMP + clip parity (arena=%s): %s" % (arena, verdict)
    # a parameter unused on rank 0 only (ref ProcessUnusedDenseVars reducer.cc:941-1011): its reduced gradient is the
    # scaled sum of the OTHER ranks' gradients on every rank
    up, ug = [], []
    for i, sz in enumerate(sizes[:7]):
        p = prb.Tensor.zeros(sz, "float32", name="u_%d" % i)
        per_rank = [np.random.default_rng(7000 * r + i).normal(0, 1e-2, sz).astype(np.float32) for r in range(n)]
        p.grad = prb.Tensor.from_numpy(per_rank[rank])
        up.append(p)
        ug.append(per_rank)
    dpu = prb.DataParallel(up, comm_buffer_size=0.4, last_comm_buffer_size=0.01, find_unused_parameters=True)
    lone = 3
    dpu.prepare_for_backward(unused_vars=[lone] if rank == 0 else [])
    for i in reversed(range(len(up))):
        if not (rank == 0 and i == lone):
            dpu.mark_var_ready(i)
    assert dpu.finalized
    prb.current_stream().synchronize()
    for i, p in enumerate(up):
        ranks = [r for r in range(n) if not (r == 0 and i == lone)]
        want = np.sum([orc.scale(ug[i][r], orc.F32, 1.0 / n).astype(np.float64) for r in ranks], axis=0)
        if n == 2:
            np.testing.assert_array_equal(p.grad.numpy(), (want + 0.0).astype(np.float32))
        else:
            np.testing.assert_allclose(p.grad.numpy(), want, rtol=1e-6, atol=1e-8)
        np.save(os.path.join(out_dir, "unused_grad_%d_r%d.npy" % (i, rank)), p.grad.numpy())
    # sparse (SelectedRows) gradients, ref EagerReducer::AllReduceSparse reducer.cc:1072-1277 and
    # adam_dense_param_sparse_grad: an embedding table next to dense parameters; the ranks hold the same number of rows
    # in step 0 (the all-gather branch of the reference) and different numbers afterwards (its broadcast branch)
    height, width = 300, 24
    emb0 = np.random.default_rng(7).normal(0, 0.02, (height, width)).astype(np.float32)
    emb = prb.Tensor.from_numpy(emb0.reshape(-1), name="emb")
    emb.is_sparse_grad = True
    dn = [prb.Tensor.from_numpy(np.random.default_rng(100 + i).normal(0, 0.02, sz).astype(np.float32), name="sd_%d" % i)
          for i, sz in enumerate(sizes[:4])]
    dps = prb.DataParallel([emb] + dn, comm_buffer_size=0.4, last_comm_buffer_size=0.01)
    opts = prb.optimizer.Adam(learning_rate=1e-3, parameters=[emb] + dn, lazy_mode=False)
    eo, m1e, m2e = emb0.copy(), np.zeros_like(emb0), np.zeros_like(emb0)
    b1p, b2p = np.float32(0.9), np.float32(0.999)
    for step in range(3):
        per_rank = []
        for r in range(n):
            rg = np.random.default_rng(5000 + 10 * step + r)
            cnt = 11 if step == 0 else 5 + 3 * r + step
            rws = rg.permutation(height)[:cnt]       # unique within a rank: a row occurs at most nranks times
            per_rank.append((rws, rg.normal(0, 1e-2, (cnt, width)).astype(np.float32)))
        dps.prepare_for_backward()
        for i, p in enumerate(dn):
            p.grad.copy_from_numpy(np.random.default_rng(6000 + 100 * step + 10 * rank + i).normal(0, 1e-2, p.numel).astype(np.float32))
        emb.grad = prb.SelectedRows(per_rank[rank][0], height, prb.Tensor.from_numpy(per_rank[rank][1].reshape(-1)))
        for i in reversed(range(len(dn) + 1)):
            dps.mark_var_ready(i)
        assert dps.finalized
        want_rows, want_vals = orc.sparse_allreduce(per_rank, orc.F32)
        np.testing.assert_array_equal(emb.grad.rows, want_rows)
        np.testing.assert_array_equal(emb.grad.value.numpy().reshape(-1, width), want_vals)
        opts.step()
        prb.current_stream().synchronize()
        orc.sparse_adam(eo, want_rows, want_vals, m1e, m2e, None, dtype_param=orc.F32, lr=1e-3, beta1=0.9, beta2=0.999,
                        epsilon=1e-8, beta1_pow=b1p, beta2_pow=b2p, lazy_mode=False)
        b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
        if n == 2:   # two addends per row at most: the merge order cannot matter
            np.testing.assert_array_equal(emb.numpy().reshape(height, width), eo)
        else:
            np.testing.assert_allclose(emb.numpy().reshape(height, width), eo, rtol=1e-5, atol=1e-7)
        np.save(os.path.join(out_dir, "sparse_emb_%d_r%d.npy" % (step, rank)), emb.numpy())
        opts.clear_grad()
    # sharded optimizer state (ZeRO-1, ref dygraph_sharding_optimizer.py): each parameter is updated by its owner only
    # and broadcast; every rank must end with exactly the bits of the replicated optimizer
    def make(prefix):
        ps = []
        for i, sz in enumerate(sizes):
            p = prb.Tensor.from_numpy(np.random.default_rng(100 + i).normal(0, 0.02, sz).astype(np.float32), name="%s_%d" % (prefix, i))
            p.grad = prb.Tensor.from_numpy(grads_all[i][rank])
            ps.append(p)
        return ps
    pa, pb = make("rep"), make("shd")
    dpa = prb.DataParallel(pa, comm_buffer_size=0.4, last_comm_buffer_size=0.01, consume_from_buckets=True)
    dpb = prb.DataParallel(pb, comm_buffer_size=0.4, last_comm_buffer_size=0.01, consume_from_buckets=True)
    oa = prb.optimizer.AdamW(learning_rate=1e-3, parameters=pa, weight_decay=0.01)
    ob = prb.distributed.DygraphShardingOptimizer(prb.distributed.ShardingGroup(group), None, pb, prb.optimizer.AdamW,
                                                  learning_rate=1e-3, weight_decay=0.01)
    owned = ob._rank2params[rank]
    assert 0 < len(owned) < len(pb) and set(ob._param2rank.values()) == set(range(n))
    for step in range(3):
        for dp_, op_ in ((dpa, oa), (dpb, ob)):
            dp_.prepare_for_backward()
            dp_.mark_all_ready()
            op_.step()
        prb.current_stream().synchronize()
        for i, (a, b) in enumerate(zip(pa, pb)):
            assert a.numpy().tobytes() == b.numpy().tobytes(), "sharded step %d param %d" % (step, i)
            np.save(os.path.join(out_dir, "shard_param_%d_%d_r%d.npy" % (step, i, rank)), b.numpy())
    assert set(ob._accumulators) == {p.name for p in owned}   # moments exist for the local shard only
    # the B200-native sharded step (GroupShardedOptimizerStage2: reduce-scatter -> update of the owned arena ranges ->
    # gather of the parameters) against the replicated optimizer: fp32, no AMP, no clip; then AMP bf16 + master + clip
    # with an Inf on one rank against the oracle
    pc = make("gs2")
    dpc = prb.DataParallel(pc, comm_buffer_size=0.4, last_comm_buffer_size=0.01, consume_from_buckets=True)
    oc = prb.distributed.GroupShardedOptimizerStage2(pc, prb.optimizer.AdamW(learning_rate=1e-3, parameters=pc, weight_decay=0.01), group)
    mine = dpc.shard_ranges()
    everyone = [dpc.shard_ranges(r) for r in range(n)]
    assert sum(ln for rr in everyone for _, _, ln in rr) == sum(sizes) and mine == everyone[rank]
    pd = make("rep2")
    dpd = prb.DataParallel(pd, comm_buffer_size=0.4, last_comm_buffer_size=0.01, consume_from_buckets=True)
    od = prb.optimizer.AdamW(learning_rate=1e-3, parameters=pd, weight_decay=0.01)
    for step in range(4):
        for dp_, op_ in ((dpd, od), (dpc, oc)):
            dp_.prepare_for_backward()
            if step % 2 == 0:
                dp_.mark_all_ready()
            else:
                for i in reversed(range(len(pc))):
                    dp_.mark_var_ready(i)
            op_.step()
        prb.current_stream().synchronize()
        assert dpc._sharded == (dpc._reducer.allreduce_backend == "nvlink")
        for i, (a, b) in enumerate(zip(pd, pc)):
            assert a.numpy().tobytes() == b.numpy().tobytes(), "GroupShardedOptimizerStage2 step %d param %d" % (step, i)
            np.save(os.path.join(out_dir, "gs2_param_%d_%d_r%d.npy" % (step, i, rank)), b.numpy())
    assert sum(v.numel for v in oc.local_params) == sum(ln for _, _, ln in mine)
    chunks = [sum((ln + 8191) // 8192 for _, _, ln in rr) for rr in everyone]   # equal shares, to within one chunk
    assert max(chunks) - min(chunks) <= 1, chunks
    for arena in (False, True):
        for layout in (True, "bucket"):   # one reduce-scatter for the whole arena / one per bucket
            verdict = bench.parity_preamble(prb, group, arena=arena, save_dir=out_dir, sharded=layout)
            assert verdict == "ok", "sharded AMP + clip parity (arena=%s, layout=%s): %s" % (arena, layout, verdict)
    # timing helper used by bench.py: max over ranks through the communicator
    t = prb.Tensor.from_numpy(np.array([float(rank + 1)], np.float32))
    group.all_reduce(t, "max")
    assert t.numpy()[0] == float(n)
    group.barrier()
    print("rank %d ok" % rank, flush=True)


def timeout_main():
    """A rank that arrives later than PRB_NVL_TIMEOUT_MS: no rank may apply a half-reduced gradient silently.  The
    waiting rank gives up, tells its peer, both exchanges leave the data alone and both hosts get an error."""
    import time
    group = prb.distributed.init_parallel_env()
    rank = group.rank
    p = prb.Tensor.from_numpy(np.full(100000, 1.0, np.float32), name="t_0")
    p.grad = prb.Tensor.from_numpy(np.full(100000, float(rank + 1), np.float32))
    dp = prb.DataParallel([p], consume_from_buckets=True)
    if dp._reducer.allreduce_backend != "nvlink":
        print("rank %d timeout skipped (backend %s)" % (rank, dp._reducer.allreduce_backend), flush=True)
        return
    dp.prepare_for_backward()
    dp.mark_all_ready()
    dp.wait_all_groups()
    prb.current_stream().synchronize()
    np.testing.assert_array_equal(dp.bucket_grads()[0].numpy(), np.full(100000, 1.5, np.float32))
    group.barrier()
    late = float(os.environ.get("PRB_TEST_LATE_S", "4"))
    if rank == 1:
        time.sleep(late)
    dp.prepare_for_backward()
    dp.mark_all_ready()
    prb.synchronize()
    time.sleep(late + 1 if rank == 0 else 0.5)  # let the peer run into it too
    raised = False
    try:
        dp.wait_all_groups()
    except prb._lib.PRBError as e:
        raised = "bucket exchange failed" in str(e)
    assert raised, "rank %d: the dead exchange was not reported" % rank
    assert dp._reducer.comm_error()
    # the bucket still holds this rank's own packed gradient: nothing half-reduced was stored
    np.testing.assert_array_equal(dp.bucket_grads()[0].numpy(), np.full(100000, (rank + 1) * 0.5, np.float32))
    print("rank %d timeout ok" % rank, flush=True)
    sys.stdout.flush()
    os._exit(0)  # the communicator is dead by design: skip the collective teardown


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "timeout":
        timeout_main()
    else:
        main(sys.argv[1])
