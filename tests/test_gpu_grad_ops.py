"""GPU parity: check_finite_and_unscale_, squared_l2_norm, ClipGradByGlobalNorm, scale, concat/split,
update_loss_scaling_ -- through the C ABI, against the oracle and the reference-derived fixtures."""
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    return prb


# ------------------------------------------------------------------ check_finite_and_unscale
@pytest.mark.parametrize("kind,expect", [("clean", False), ("nan", True), ("inf", True)])
def test_check_finite_reference_optest(prb, golden_dir, kind, expect):
    """test_amp_check_finite_and_scale_op.py:29-90: x (1024,1024) random from the reference's seed."""
    fx = np.load(os.path.join(golden_dir, "check_finite_optest.npz"))
    np.random.seed(123)
    x = np.random.random((1024, 1024)).astype(np.float32)
    scale = np.random.random((1)).astype(np.float32)
    if kind != "clean":
        x[128][128] = np.nan if kind == "nan" else np.inf
    t = prb.Tensor.from_numpy(x.ravel())
    s = prb.Tensor.from_numpy(scale)
    _, found = prb._C_ops.check_finite_and_unscale_([t], s)
    assert bool(found.numpy()[0]) == expect == bool(fx[f"found_{kind}"][0])
    ref = x.ravel().copy()
    assert orc.check_finite_and_unscale([ref], orc.F32, float(scale[0])) == expect
    if kind == "clean":
        bits_equal(t.numpy(), ref)
        np.testing.assert_allclose(t.numpy().reshape(1024, 1024)[::16, ::16], fx["out_sample"], rtol=2e-7, atol=1e-5)


@pytest.mark.parametrize("dtype", [orc.F32, orc.BF16, orc.F16])
def test_check_finite_ragged_empty_and_multi_tensor(prb, dtype):
    rng = np.random.default_rng(3)
    sizes = [10, 0, 10, 8192, 8193, 1, 0, 70000]
    xs32 = [(rng.normal(0, 1, n) * 2.0 ** 15 * 1e-2).astype(np.float32) for n in sizes]
    xs = [orc.from_f32(x, dtype) for x in xs32]
    dev = [prb.Tensor.from_numpy(x, dtype) for x in xs]
    s = prb.Tensor.from_numpy(np.array([2.0 ** 15], np.float32))
    _, found = prb._C_ops.check_finite_and_unscale_(dev, s)
    ref = [x.copy() for x in xs]
    f = orc.check_finite_and_unscale(ref, dtype, 2.0 ** 15)
    assert bool(found.numpy()[0]) == f == False  # noqa: E712
    for d, r in zip(dev, ref):
        bits_equal(d.numpy(), r)
    # one poisoned element in the last tensor flips the flag; the flag is reset on every call
    xs[-1][12345] = orc.from_f32(np.array([np.inf], np.float32), dtype)[0]
    dev[-1].copy_from_numpy(xs[-1])
    _, found = prb._C_ops.check_finite_and_unscale_(dev, s, found)
    assert bool(found.numpy()[0])
    dev[-1].copy_from_numpy(ref[-1])
    _, found = prb._C_ops.check_finite_and_unscale_(dev, s, found)
    assert not bool(found.numpy()[0])


def test_check_finite_on_the_narrowed_value(prb):
    x = prb.Tensor.from_numpy(np.array([0x7F7F], np.uint16), "bfloat16")  # bf16 max: finite before, inf after narrowing
    s = prb.Tensor.from_numpy(np.array([0.997], np.float32))
    _, found = prb._C_ops.check_finite_and_unscale_([x], s)
    assert bool(found.numpy()[0]) and x.numpy()[0] == 0x7F80
    _, found = prb._C_ops.check_finite_and_unscale_([], s)  # empty list: flag cleared, nothing else (amp_kernel.cu:283)
    assert not bool(found.numpy()[0])


@pytest.mark.parametrize("dtype", [orc.BF16, orc.F16])
def test_16bit_packed_pairs_edge_values(prb, dtype):
    """The 16-bit pass works on packed pairs: subnormal products, values that overflow when narrowed, NaN/Inf in
    either half of a pair, odd lengths -- all must agree with the element-wise oracle bit for bit."""
    rng = np.random.default_rng(21)
    n = 8192 * 2 + 5
    mag = 10.0 ** rng.uniform(-30 if dtype == orc.BF16 else -7, 3, n)
    x32 = (mag * rng.choice([-1.0, 1.0], n)).astype(np.float32)
    x = orc.from_f32(x32, dtype)
    t = prb.Tensor.from_numpy(x, dtype)
    s = prb.Tensor.from_numpy(np.array([0.37], np.float32))
    _, found = prb._C_ops.check_finite_and_unscale_([t], s)
    ref = x.copy()
    f = orc.check_finite_and_unscale([ref], dtype, 0.37)
    bits_equal(t.numpy(), ref)
    assert bool(found.numpy()[0]) == f
    got = orc.to_f32(prb._C_ops.squared_l2_norm(t).numpy(), dtype)[0]
    _, want64 = orc.squared_l2_norm(ref, dtype)
    if np.isfinite(want64) and want64 < (6e4 if dtype == orc.F16 else 3e38):
        np.testing.assert_allclose(got, want64, rtol=8e-3 if dtype == orc.BF16 else 2e-3)
    else:
        assert not np.isfinite(got)
    for pos in (0, 1, n - 1, 8191, 8192):  # a poisoned element in the low half, the high half, the ragged tail
        for poison in (np.inf, np.nan):
            y = x.copy()
            y[pos] = orc.from_f32(np.array([poison], np.float32), dtype)[0]
            t.copy_from_numpy(y)
            _, found = prb._C_ops.check_finite_and_unscale_([t], s, found)
            assert bool(found.numpy()[0]), (pos, poison)
    small = orc.from_f32(np.array([3e-5, -2e-5, 1e-6, 0.0, -0.0, 7e-5, 6e-8, 1.0], np.float32), dtype)
    ts = prb.Tensor.from_numpy(small, dtype)
    got = prb._C_ops.squared_l2_norm(ts).numpy()
    want, _ = orc.squared_l2_norm(small, dtype)   # sequential oracle: same order for 8 elements of one thread
    bits_equal(got, orc.from_f32(np.array([want], np.float32), dtype))


# ------------------------------------------------------------------ squared_l2_norm / clip
def test_squared_l2_norm_reference_cases(prb, golden_dir):
    fx = np.load(os.path.join(golden_dir, "squared_l2_norm_optest.npz"))
    X = fx["TestL2LossOp/X"].ravel()
    out = prb._C_ops.squared_l2_norm(prb.Tensor.from_numpy(X)).numpy()[0]
    np.testing.assert_allclose(out, fx["TestL2LossOp/Out"][0], rtol=1e-5)
    for name in ("F16Op", "F16Op1", "F16Op2"):
        x32 = fx[f"{name}/X"].ravel()
        v32 = prb._C_ops.squared_l2_norm(prb.Tensor.from_numpy(x32)).numpy()[0]
        x16 = orc.from_f32(x32, orc.F16)
        v16 = orc.to_f32(prb._C_ops.squared_l2_norm(prb.Tensor.from_numpy(x16, "float16")).numpy(), orc.F16)[0]
        want16, _ = orc.squared_l2_norm(x16, orc.F16)
        np.testing.assert_allclose(v16, want16, rtol=4e-3)  # T-typed output: 1 ulp of fp16 ~ 1e-3
        np.testing.assert_allclose(v16, v32, rtol=4e-3)


@pytest.mark.parametrize("dtype,n", [(orc.F32, 1), (orc.F32, 5 * 11 * 13), (orc.F32, (1 << 22) + 5), (orc.BF16, 300001),
                                    (orc.F16, 8192 * 3)])
def test_squared_l2_norm_vs_fp64_and_deterministic(prb, dtype, n):
    rng = np.random.default_rng(n)
    x = orc.from_f32(rng.normal(0, 1e-2, n).astype(np.float32), dtype)
    t = prb.Tensor.from_numpy(x, dtype)
    y1 = prb._C_ops.squared_l2_norm(t).numpy()
    y2 = prb._C_ops.squared_l2_norm(t).numpy()
    bits_equal(y1, y2)  # run-to-run determinism (test_squared_l2_norm_op.py:98-111)
    _, want64 = orc.squared_l2_norm(x, dtype)
    got = orc.to_f32(y1, dtype)[0]
    tol = 1e-5 if dtype == orc.F32 else (8e-3 if dtype == orc.BF16 else 1e-3)  # summation order + T-typed output
    np.testing.assert_allclose(got, want64, rtol=tol)


@pytest.mark.parametrize("dtype", [orc.F32, orc.BF16])
@pytest.mark.parametrize("auto_skip", [False, True])
@pytest.mark.parametrize("gscale", [1.0, 1e-3])
def test_clip_grad_by_global_norm(prb, dtype, auto_skip, gscale):
    """python/paddle/nn/clip.py:643-736 and test_gradient_clip.py:442-482 (norm after clip == min(norm, clip))."""
    rng = np.random.default_rng(7)
    sizes = [64, 3 * 3 * 64 * 64, 1000, 2048, 1, 70001]
    gs = [orc.from_f32((rng.normal(0, 1e-2, n) * gscale).astype(np.float32), dtype) for n in sizes]
    dev = [prb.Tensor.from_numpy(g, dtype) for g in gs]
    norm_t, clip_t = prb._C_ops.clip_grad_by_global_norm_(dev, 1.0, auto_skip)
    ref = [g.copy() for g in gs]
    norm, clip = orc.global_norm_clip(ref, [dtype] * len(ref), 1.0, auto_skip)
    tol = 2e-6 if dtype == orc.F32 else 4e-3
    np.testing.assert_allclose(norm_t.numpy()[0], norm, rtol=tol)
    got_clip = clip_t.numpy()[0]
    if clip is None:
        assert got_clip == 1.0
        for d, g in zip(dev, gs):
            bits_equal(d.numpy(), g)
    else:
        np.testing.assert_allclose(got_clip, clip, rtol=tol)
        # the multiply is exact given the clip factor: re-run the oracle's multiply with the device factor
        for d, g in zip(dev, gs):
            r = g.copy()
            orc.clip_mul(r, dtype, float(got_clip))
            bits_equal(d.numpy(), r)
        if dtype == orc.F32:
            after = np.sqrt(sum(float(np.sum(d.numpy().astype(np.float64) ** 2)) for d in dev))
            np.testing.assert_allclose(after, min(norm, 1.0), rtol=1e-5)


def test_clip_mixed_dtypes_one_call(prb):
    rng = np.random.default_rng(8)
    g32 = rng.normal(0, 1, 5000).astype(np.float32)
    g16 = orc.from_f32(rng.normal(0, 1, 7000).astype(np.float32), orc.BF16)
    dev = [prb.Tensor.from_numpy(g32), prb.Tensor.from_numpy(g16, "bfloat16")]
    norm_t, clip_t = prb._C_ops.clip_grad_by_global_norm_(dev, 1.0)
    ref = [g32.copy(), g16.copy()]
    norm, clip = orc.global_norm_clip(ref, [orc.F32, orc.BF16], 1.0)
    np.testing.assert_allclose(norm_t.numpy()[0], norm, rtol=4e-3)
    np.testing.assert_allclose(clip_t.numpy()[0], clip, rtol=4e-3)


# ------------------------------------------------------------------ bucket ops
def test_scale_semantics(prb):
    x = np.array([-0.0, 1.0, -3.0, 1e-45, 7.0], np.float32)
    for nranks in (1, 2, 3, 8):
        out = prb._C_ops.scale(prb.Tensor.from_numpy(x), 1.0 / nranks, 0.0, False).numpy()
        bits_equal(out, orc.scale(x, orc.F32, 1.0 / nranks))
    assert not np.signbit(prb._C_ops.scale(prb.Tensor.from_numpy(x), 1.0, 0.0, False).numpy()[0])  # -0 + 0 -> +0
    out = prb._C_ops.scale(prb.Tensor.from_numpy(x), 0.5, 1.0, True).numpy()
    bits_equal(out, orc.scale(x, orc.F32, 0.5, 1.0, True))
    rng = np.random.default_rng(0)
    for dt in (orc.BF16, orc.F16):
        xb = orc.from_f32(rng.normal(0, 1, 100003).astype(np.float32), dt)
        t = prb.Tensor.from_numpy(xb, dt)
        prb._C_ops.scale_(t, 1.0 / 8, 0.0, False)
        bits_equal(t.numpy(), orc.scale(xb, dt, 1.0 / 8))
    empty = prb.Tensor(0, "float32")
    prb._C_ops.scale_(empty, 0.5)  # numel <= 0 returns (scale_kernel.cu:59-61)


@pytest.mark.parametrize("np_dtype,dtype", [(np.float32, "float32"), (np.uint16, "bfloat16"), (np.float64, "float64")])
def test_concat_split_bit_exact_dense_and_ragged(prb, np_dtype, dtype):
    """Dense reference packing: segment starts are only element-aligned, sizes ragged, some empty."""
    rng = np.random.default_rng(5)
    sizes = [1, 7, 0, 8, 8191, 8192, 8193, 3, 100000, 0, 5]
    if np_dtype == np.uint16:
        arrs = [rng.integers(0, 65536, n).astype(np.uint16) for n in sizes]
    else:
        arrs = [rng.normal(0, 1, n).astype(np_dtype) for n in sizes]
        arrs[3][:2] = [-0.0, np.nan]
    dev = [prb.Tensor.from_numpy(a, dtype) for a in arrs]
    bucket = prb._C_ops.concat(dev)
    bits_equal(bucket.numpy(), orc.concat(arrs) if sum(sizes) else np.array([], np_dtype))
    outs = [prb.Tensor.zeros(n, dtype) for n in sizes]
    prb._C_ops.split_(bucket, outs)
    for o, a in zip(outs, arrs):
        bits_equal(o.numpy(), a)


def test_concat_rejects_mixed_dtypes_and_wrong_sizes(prb):
    from paddle_ray_b200._lib import PRBError
    a, b = prb.Tensor.zeros(4, "float32"), prb.Tensor.zeros(4, "bfloat16")
    with pytest.raises(PRBError, match="unexpected dtype"):
        prb._C_ops.concat([a, b])
    with pytest.raises(PRBError, match="do not add up"):
        prb._C_ops.split_(prb.Tensor.zeros(9, "float32"), [prb.Tensor.zeros(4, "float32"), prb.Tensor.zeros(4, "float32")])


# ------------------------------------------------------------------ update_loss_scaling
def test_update_loss_scaling_matches_oracle_state_machine(prb):
    T = prb.Tensor
    scale = T.from_numpy(np.array([2.0 ** 15], np.float32))
    good, bad = T.zeros(1, "int32"), T.zeros(1, "int32")
    found = T.zeros(1, "bool")
    x = T.from_numpy(np.arange(1, 10001, dtype=np.float32))
    ls, g, b = 2.0 ** 15, 0, 0
    pattern = [False] * 5 + [True, True] + [False] * 3 + [True] + [False] * 4
    for f in pattern:
        found.copy_from_numpy(np.array([f]))
        prb._C_ops.update_loss_scaling_([x], found, scale, good, bad, 3, 2, 2.0, 0.5)
        ls, g, b = orc.update_loss_scaling(f, ls, g, b, 3, 2, 2.0, 0.5)
        assert (scale.numpy()[0], good.numpy()[0], bad.numpy()[0]) == (ls, g, b)
        if f:  # LazyZeros: gradients are zeroed on a bad step (amp_kernel_impl.h:127-136)
            assert not x.numpy().any()
            x.copy_from_numpy(np.arange(1, 10001, dtype=np.float32))
        else:
            assert x.numpy()[-1] == 10000.0
