"""adam_ with a SelectedRows gradient (adam_dense_param_sparse_grad, selected_rows/gpu/adam_kernel.cu:95-260) and the
reducer's sparse groups (EagerReducer::AllReduceSparse, reducer.cc:1072-1277), through the C ABI, against the oracle
and the reference's own TestSparseAdamOp fixture."""
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import bits_equal

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    return prb


def _run_op(prb, p0, rows, vals, m10, m20, master0, dtype, lazy, b1p0, b2p0, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8,
            skip=None, use_global_beta_pow=False, pows_on_host=False, skip_on_host=False):
    T = prb.Tensor
    height = p0.shape[0]
    p = T.from_numpy(p0.reshape(-1), dtype)
    m1, m2 = T.from_numpy(m10.reshape(-1)), T.from_numpy(m20.reshape(-1))
    master = T.from_numpy(master0.reshape(-1)) if master0 is not None else None
    value = T.from_numpy(np.ascontiguousarray(vals).reshape(-1), dtype) if vals.size else T(0, dtype)
    g = prb.SelectedRows(rows, height, value)
    lrt = T.from_numpy(np.array([lr], np.float32))
    dev = -1 if pows_on_host else None
    b1p, b2p = T.from_numpy(np.array([b1p0], np.float32), device=dev), T.from_numpy(np.array([b2p0], np.float32), device=dev)
    sk = T.from_numpy(np.array([skip], np.bool_), device=-1 if skip_on_host else None) if skip is not None else None
    prb._C_ops.adam_(p, g, lrt, m1, m2, b1p, b2p, master, sk, beta1, beta2, eps, lazy, 1000, master is not None,
                     use_global_beta_pow)
    shape = p0.shape
    return (p.numpy().reshape(shape), m1.numpy().reshape(shape), m2.numpy().reshape(shape),
            None if master is None else master.numpy().reshape(shape), b1p.numpy()[0], b2p.numpy()[0])


def test_reference_optest_fixture(prb):
    """TestSparseAdamOp (test_adam_op.py:355-460): the reference's own expected outputs, its own acceptance criterion,
    and bit equality with the oracle."""
    z = np.load(os.path.join(GOLDEN, "sparse_adam_optest.npz"))
    beta1, beta2, eps = (float(v) for v in z["hyper"])
    rows = z["rows"]
    lr = float(z["in/LearningRate"][0])
    for tag, lazy in (("full", False), ("lazy", True)):
        p0, m10, m20 = (z["in/" + k].copy() for k in ("Param", "Moment1", "Moment2"))
        got = _run_op(prb, p0, rows, z["grad"], m10, m20, None, orc.F32, lazy, z["in/Beta1Pow"][0], z["in/Beta2Pow"][0],
                      lr=lr, beta1=beta1, beta2=beta2, eps=eps)
        po, m1o, m2o = p0.copy(), m10.copy(), m20.copy()
        orc.sparse_adam(po, rows, z["grad"], m1o, m2o, None, dtype_param=orc.F32, lr=lr, beta1=beta1, beta2=beta2,
                        epsilon=eps, beta1_pow=z["in/Beta1Pow"][0], beta2_pow=z["in/Beta2Pow"][0], lazy_mode=lazy)
        for a, b in zip(got[:3], (po, m1o, m2o)):
            bits_equal(a, b)
        touched = rows if lazy else np.arange(p0.shape[0])
        for a, key in zip(got[:3], ("/ParamOut", "/Moment1Out", "/Moment2Out")):
            want = z[tag + key][touched]
            assert ((a[touched] - want) < 1e-5).all(), key                   # the reference's criterion (:442-443)
            np.testing.assert_allclose(a[touched], want, atol=2e-5, rtol=0)  # two-sided (epsilon factoring, see oracle test)
        assert got[4] == np.float32(beta1) * z["in/Beta1Pow"][0] and got[5] == np.float32(beta2) * z["in/Beta2Pow"][0]


@pytest.mark.parametrize("lazy", [False, True])
@pytest.mark.parametrize("dtype,master", [(orc.F32, False), (orc.BF16, True), (orc.F16, True), (orc.BF16, False)])
def test_random_rows_with_duplicates_bit_exact(prb, dtype, master, lazy):
    rng = np.random.default_rng(17)
    for height, width, nrows in ((50, 8, 20), (1000, 64, 300), (30000, 128, 4096), (7, 5, 40), (64, 768, 1)):
        rows = rng.integers(0, height, nrows)            # unsorted, with duplicates (up to many per row)
        p32 = rng.normal(0, 0.02, (height, width)).astype(np.float32)
        p0 = orc.from_f32(p32.reshape(-1), dtype).reshape(height, width)
        vals = orc.from_f32(rng.normal(0, 1e-2, nrows * width).astype(np.float32), dtype).reshape(nrows, width)
        m10 = rng.normal(0, 1e-3, (height, width)).astype(np.float32)
        m20 = rng.uniform(0, 1e-3, (height, width)).astype(np.float32)
        ms0 = orc.to_f32(p0.reshape(-1), dtype).reshape(height, width) if master else None
        b1p, b2p = np.float32(0.9 ** 3), np.float32(0.999 ** 3)
        got = _run_op(prb, p0, rows, vals, m10, m20, ms0, dtype, lazy, b1p, b2p)
        po, m1o, m2o = p0.copy(), m10.copy(), m20.copy()
        mso = None if ms0 is None else ms0.copy()
        orc.sparse_adam(po, rows, vals, m1o, m2o, mso, dtype_param=dtype, lr=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8,
                        beta1_pow=b1p, beta2_pow=b2p, lazy_mode=lazy)
        bits_equal(got[0], po)
        bits_equal(got[1], m1o)
        bits_equal(got[2], m2o)
        if master:
            bits_equal(got[3], mso)
        if lazy:  # rows without a gradient keep their bits
            absent = np.setdiff1d(np.arange(height), rows)
            bits_equal(got[0][absent], p0[absent])
            bits_equal(got[1][absent], m10[absent])


def test_sorted_unique_rows_use_the_values_in_place_and_edge_cases(prb):
    rng = np.random.default_rng(3)
    height, width = 40, 16
    p0 = rng.normal(0, 0.02, (height, width)).astype(np.float32)
    m10 = rng.normal(0, 1e-3, (height, width)).astype(np.float32)
    m20 = rng.uniform(0, 1e-3, (height, width)).astype(np.float32)
    b1p, b2p = np.float32(0.9), np.float32(0.999)
    for rows in (np.array([0, 3, 39]), np.arange(height), np.array([], np.int64)):
        vals = rng.normal(0, 1e-2, (rows.size, width)).astype(np.float32)
        for lazy in (False, True):
            got = _run_op(prb, p0, rows, vals, m10, m20, None, orc.F32, lazy, b1p, b2p)
            po, m1o, m2o = p0.copy(), m10.copy(), m20.copy()
            orc.sparse_adam(po, rows, vals, m1o, m2o, None, dtype_param=orc.F32, lr=1e-3, beta1=0.9, beta2=0.999,
                            epsilon=1e-8, beta1_pow=b1p, beta2_pow=b2p, lazy_mode=lazy)
            bits_equal(got[0], po)
            bits_equal(got[1], m1o)
            bits_equal(got[2], m2o)
    # skip_update: nothing moves, the accumulators stay; use_global_beta_pow: the accumulators stay
    rows = np.array([5, 1, 5])
    vals = rng.normal(0, 1e-2, (3, width)).astype(np.float32)
    got = _run_op(prb, p0, rows, vals, m10, m20, None, orc.F32, False, b1p, b2p, skip=True)
    bits_equal(got[0], p0)
    bits_equal(got[1], m10)
    assert got[4] == b1p and got[5] == b2p
    got = _run_op(prb, p0, rows, vals, m10, m20, None, orc.F32, False, b1p, b2p, use_global_beta_pow=True)
    assert got[4] == b1p and got[5] == b2p and not np.array_equal(got[0], p0)
    # beta-pow accumulators on the host (the reference's default placement): same bits, advanced on the host; a skip
    # flag on the device is read back first, like the reference does
    dev_run = _run_op(prb, p0, rows, vals, m10, m20, None, orc.F32, True, b1p, b2p)
    host_run = _run_op(prb, p0, rows, vals, m10, m20, None, orc.F32, True, b1p, b2p, pows_on_host=True)
    for a, b in zip(dev_run[:3], host_run[:3]):
        bits_equal(a, b)
    assert host_run[4] == dev_run[4] == np.float32(0.9) * b1p and host_run[5] == dev_run[5]
    for skip_on_host in (False, True):
        got = _run_op(prb, p0, rows, vals, m10, m20, None, orc.F32, True, b1p, b2p, pows_on_host=True, skip=True,
                      skip_on_host=skip_on_host)
        bits_equal(got[0], p0)
        assert got[4] == b1p and got[5] == b2p
    # a row outside the table is refused before anything is launched
    with pytest.raises(prb._lib.PRBError):
        _run_op(prb, p0, np.array([height]), vals[:1], m10, m20, None, orc.F32, False, b1p, b2p)


def test_adam_optimizer_with_an_embedding_table_and_dense_parameters(prb):
    """Optimizer.step with one sparse and several dense parameters through DataParallel (one rank: the sparse group's
    exchange is the 1/nranks scale alone): three steps, bit-exact against the oracle's per-parameter composition; the
    shared beta-pow accumulators advance once per step."""
    T = prb.Tensor
    rng = np.random.default_rng(23)
    height, width = 500, 32
    sizes = [1000, 4097, 64]
    emb0 = rng.normal(0, 0.02, (height, width)).astype(np.float32)
    dense0 = [rng.normal(0, 0.02, n).astype(np.float32) for n in sizes]
    emb = T.from_numpy(emb0.reshape(-1), name="emb")
    emb.is_sparse_grad = True
    dense = [T.from_numpy(a, name="w_%d" % i) for i, a in enumerate(dense0)]
    params = [emb] + dense
    group = prb.distributed.init_parallel_env()
    dp = prb.DataParallel(params, group=group)
    assert [0] in dp.group_indices            # the sparse parameter is a group of its own (reducer.cc:94-99)
    opt = prb.optimizer.Adam(learning_rate=1e-3, parameters=params, lazy_mode=True)
    eo, m1e, m2e = emb0.copy(), np.zeros_like(emb0), np.zeros_like(emb0)
    do = [a.copy() for a in dense0]
    m1d, m2d = [np.zeros_like(a) for a in dense0], [np.zeros_like(a) for a in dense0]
    b1p, b2p = np.float32(0.9), np.float32(0.999)
    for step in range(3):
        rows = rng.integers(0, height, 37)
        vals = rng.normal(0, 1e-2, (rows.size, width)).astype(np.float32)
        gd = [rng.normal(0, 1e-2, n).astype(np.float32) for n in sizes]
        dp.prepare_for_backward()
        for p, g in zip(dense, gd):
            p.grad.copy_from_numpy(g)
        emb.grad = prb.SelectedRows(rows, height, T.from_numpy(vals.reshape(-1)))
        dp.mark_all_ready()
        assert dp.finalized
        opt.step()
        opt.clear_grad()
        assert emb.grad is None
        kw = dict(lr=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8, beta1_pow=b1p, beta2_pow=b2p)
        r2, v2 = orc.sparse_allreduce([(rows, vals)], orc.F32)
        orc.sparse_adam(eo, r2, v2, m1e, m2e, None, dtype_param=orc.F32, lazy_mode=True, **kw)
        for i in range(len(sizes)):
            orc.adam(do[i], orc.scale(gd[i], orc.F32, 1.0), m1d[i], m2d[i], None, dtype_param=orc.F32, dtype_grad=orc.F32, **kw)
        b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
        bits_equal(emb.numpy().reshape(height, width), eo)
        for p, w in zip(dense, do):
            bits_equal(p.numpy(), w)
        assert opt._beta1_pow.numpy()[0] == b1p and opt._beta2_pow.numpy()[0] == b2p
    # AdamW has no SelectedRows kernel in the reference either
    emb.grad = prb.SelectedRows([1], height, T.from_numpy(np.ones(width, np.float32)))
    with pytest.raises(prb._lib.PRBError):
        prb.optimizer.AdamW(learning_rate=1e-3, parameters=[emb]).step()


@pytest.mark.parametrize("dtype", [orc.F32, orc.BF16, orc.F16])
def test_merge_selected_rows_and_multiply_by_clip(prb, dtype):
    """merge_selected_rows (legacy_ops.yaml:1082-1088) against the oracle's MergeAdd, and the clip's multiply
    (clip.py:715-727: the factor is first cast to the gradient's dtype) against the oracle's clip_mul."""
    T = prb.Tensor
    rng = np.random.default_rng(41)
    for height, width, nrows in ((100, 8, 37), (30522, 768, 500), (9, 5, 30), (50, 16, 0)):
        rows = rng.integers(0, height, nrows)
        vals = orc.from_f32(rng.normal(0, 1e-2, nrows * width).astype(np.float32), dtype).reshape(nrows, width)
        sr = prb.SelectedRows(rows, height, T.from_numpy(vals.reshape(-1), dtype) if nrows else T(0, dtype))
        m = prb._C_ops.merge_selected_rows(sr)
        ur, mv = orc.merge_rows(rows, vals, dtype, width)
        np.testing.assert_array_equal(m.rows, ur)
        assert m.height == height and m.value.numel == ur.size * width
        if ur.size:
            bits_equal(m.value.numpy().reshape(ur.size, width), mv)
    # already strictly ascending: the values come back unchanged
    vals = orc.from_f32(rng.normal(0, 1, 3 * 4).astype(np.float32), dtype)
    m = prb._C_ops.merge_selected_rows(prb.SelectedRows([1, 5, 7], 10, T.from_numpy(vals, dtype)))
    assert m.rows.tolist() == [1, 5, 7]
    bits_equal(m.value.numpy(), vals)
    # x = T(MT(x) * MT(T(clip)))
    x0 = orc.from_f32(rng.normal(0, 1, 10007).astype(np.float32), dtype)
    for c in (0.123456789, 1.0, 0.75):
        x = T.from_numpy(x0.copy(), dtype)
        prb._C_ops.multiply_by_clip_([x], T.from_numpy(np.array([c], np.float32)))
        want = x0.copy()
        orc.clip_mul(want, dtype, float(np.float32(c)))
        bits_equal(x.numpy(), want)


def test_global_norm_clip_spans_dense_and_sparse_gradients(prb):
    """ClipGradByGlobalNorm with a SelectedRows gradient (clip.py:643-736): the sparse gradient enters the global norm
    through its MERGED rows and is scaled as it is; Adam then consumes the clipped gradients.  Three steps against the
    oracle's composition (the clip factor within the tolerance of an unpinned summation order, everything after it
    bit for bit with that factor)."""
    T = prb.Tensor
    rng = np.random.default_rng(29)
    height, width = 200, 16
    sizes = [3000, 8193]
    emb0 = rng.normal(0, 0.02, (height, width)).astype(np.float32)
    dense0 = [rng.normal(0, 0.02, n).astype(np.float32) for n in sizes]
    emb = T.from_numpy(emb0.reshape(-1), name="emb")
    emb.is_sparse_grad = True
    dense = [T.from_numpy(a, name="w_%d" % i) for i, a in enumerate(dense0)]
    params = [emb] + dense
    clip = prb.nn.ClipGradByGlobalNorm(0.05)
    opt = prb.optimizer.Adam(learning_rate=1e-3, parameters=params, grad_clip=clip, lazy_mode=False)
    eo, m1e, m2e = emb0.copy(), np.zeros_like(emb0), np.zeros_like(emb0)
    do = [a.copy() for a in dense0]
    m1d, m2d = [np.zeros_like(a) for a in dense0], [np.zeros_like(a) for a in dense0]
    b1p, b2p = np.float32(0.9), np.float32(0.999)
    for step in range(3):
        rows = rng.integers(0, height, 50)                       # with duplicates: merged norm != raw norm
        vals = rng.normal(0, 1e-2, (rows.size, width)).astype(np.float32)
        gd = [rng.normal(0, 1e-2, n).astype(np.float32) for n in sizes]
        for p, g in zip(dense, gd):
            p.grad = T.from_numpy(g)
        emb.grad = prb.SelectedRows(rows, height, T.from_numpy(vals.reshape(-1)))
        opt.step()
        _, merged = orc.merge_rows(rows, vals, orc.F32, width)
        work = [merged.reshape(-1).copy()] + [g.copy() for g in gd]
        norm_or, c_or = orc.global_norm_clip(work, [orc.F32] * len(work), 0.05)
        c = float(clip.last_clip.numpy()[0])
        assert c < 1.0 and abs(c - c_or) <= 2e-6 * abs(c_or), (c, c_or)
        np.testing.assert_allclose(clip.last_global_norm.numpy()[0], norm_or, rtol=2e-6)
        raw_norm = np.sqrt(sum(float(np.sum(a.astype(np.float64) ** 2)) for a in [vals] + gd))
        assert abs(raw_norm - norm_or) > 1e-4 * norm_or          # the unmerged rows would give another norm
        vc = vals.copy().reshape(-1)
        orc.clip_mul(vc, orc.F32, c)
        kw = dict(lr=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8, beta1_pow=b1p, beta2_pow=b2p)
        orc.sparse_adam(eo, rows, vc.reshape(rows.size, width), m1e, m2e, None, dtype_param=orc.F32, lazy_mode=False, **kw)
        for i in range(len(sizes)):
            g = gd[i].copy()
            orc.clip_mul(g, orc.F32, c)
            orc.adam(do[i], g, m1d[i], m2d[i], None, dtype_param=orc.F32, dtype_grad=orc.F32, **kw)
        b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
        bits_equal(emb.numpy().reshape(height, width), eo)
        for p, w in zip(dense, do):
            bits_equal(p.numpy(), w)
        assert opt._beta1_pow.numpy()[0] == b1p
    # AMP on a sparse gradient: refused, as in the reference (no SelectedRows kernel for check_finite_and_unscale)
    emb.grad = prb.SelectedRows([1], height, T.from_numpy(np.ones(width, np.float32)))
    scaler = prb.amp.GradScaler()
    with pytest.raises(prb._lib.PRBError):
        scaler.step(opt)
