"""float64 operator entry points (the reference registers `double` for every op on the path: adamw_kernel.cu:347-352,
amp_kernel.cu:352-368, scale_kernel.cu:70-75, squared_l2_norm_kernel.cu:38-43) and AmpScaler.scale()."""
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


def _state(rng, n):
    return (rng.normal(0, 0.02, n), rng.normal(0, 1e-2, n), rng.normal(0, 1e-3, n), rng.uniform(0, 1e-3, n))


@pytest.mark.parametrize("pows_on_host", [False, True])
@pytest.mark.parametrize("op", ["adamw", "adam"])
def test_adam_family_float64_is_bit_exact(prb, op, pows_on_host):
    T = prb.Tensor
    rng = np.random.default_rng(5)
    for n in (1, 1000, 4097, 300001):
        p0, g0, m10, m20 = _state(rng, n)
        p, g, m1, m2 = (T.from_numpy(a.copy()) for a in (p0, g0, m10, m20))
        lr = T.from_numpy(np.array([1e-3]))
        dev = -1 if pows_on_host else None
        b1p = T.from_numpy(np.array([0.9 ** 10]), device=dev)
        b2p = T.from_numpy(np.array([0.999 ** 10]), device=dev)
        if op == "adamw":
            prb._C_ops.adamw_(p, g, lr, m1, m2, b1p, b2p, None, None, 0.9, 0.999, 1e-8, 0.5, 0.01, True)
        else:
            prb._C_ops.adam_(p, g, lr, m1, m2, b1p, b2p, None, None, 0.9, 0.999, 1e-8)
        po, m1o, m2o = p0.copy(), m10.copy(), m20.copy()
        # float attributes (lr_ratio, coeff) arrive as float32 through the ABI, like the reference's float attrs
        orc.adam(po, g0, m1o, m2o, None, dtype_param=orc.F64, dtype_grad=orc.F64, lr=1e-3,
                 lr_ratio=float(np.float32(0.5)) if op == "adamw" else 1.0, beta1=0.9, beta2=0.999, epsilon=1e-8,
                 coeff=float(np.float32(0.01)) if op == "adamw" else 0.0, use_decay=op == "adamw",
                 beta1_pow=0.9 ** 10, beta2_pow=0.999 ** 10)
        bits_equal(p.numpy(), po)
        bits_equal(m1.numpy(), m1o)
        bits_equal(m2.numpy(), m2o)
        assert b1p.numpy()[0] == 0.9 * 0.9 ** 10 and b2p.numpy()[0] == 0.999 * 0.999 ** 10


def test_merged_and_fused_adam_float64_equal_per_tensor(prb):
    T = prb.Tensor
    rng = np.random.default_rng(6)
    sizes = [7, 1000, 0, 4097]
    st = [_state(rng, n) for n in sizes]

    def fresh():
        return [[T.from_numpy(a.copy()) for a in s] for s in st]

    def pows():
        return [T.from_numpy(np.array([0.9])) for _ in sizes], [T.from_numpy(np.array([0.999])) for _ in sizes]
    lr = T.from_numpy(np.array([1e-3]))
    a = fresh()
    b1, b2 = pows()
    for (p, g, m1, m2), x, y in zip(a, b1, b2):
        if p.numel:
            prb._C_ops.adam_(p, g, lr, m1, m2, x, y, None, None, 0.9, 0.999, 1e-8)
    b = fresh()
    b1, b2 = pows()
    prb._C_ops.merged_adam_([s[0] for s in b], [s[1] for s in b], [lr] * len(sizes), [s[2] for s in b], [s[3] for s in b],
                            b1, b2, None, 0.9, 0.999, 1e-8)
    c = fresh()
    b1, b2 = pows()
    prb._C_ops.fused_adam_([s[0] for s in c], [s[1] for s in c], lr, [s[2] for s in c], [s[3] for s in c], b1, b2, None,
                           None, 0.9, 0.999, 1e-8, 4096, 0.0, False)
    for sa, sb, sc in zip(a, b, c):
        for k in (0, 2, 3):
            bits_equal(sa[k].numpy(), sb[k].numpy())
            bits_equal(sa[k].numpy(), sc[k].numpy())


def test_scale_unscale_and_squared_norm_float64(prb):
    T = prb.Tensor
    rng = np.random.default_rng(8)
    x = rng.normal(0, 3.0, 100003)
    x[5] = -0.0
    t = T.from_numpy(x)
    bits_equal(prb._C_ops.scale(t, 1.0 / 3.0, 0.0, False).numpy(), orc.scale(x, orc.F64, 1.0 / 3.0))
    bits_equal(prb._C_ops.scale(t, 0.25, 1.5, True).numpy(), orc.scale(x, orc.F64, 0.25, 1.5, True))
    # squared_l2_norm: deterministic, within rounding of the fp64 sum (order is unpinned by the reference)
    a = prb._C_ops.squared_l2_norm(t).numpy()
    b = prb._C_ops.squared_l2_norm(t).numpy()
    assert a.tobytes() == b.tobytes()
    np.testing.assert_allclose(a[0], float(np.sum(x * x)), rtol=1e-12)
    # check_finite_and_unscale: out = x * (1.0 / scale) in double, flag semantics of the reference's OpTests
    scale = T.from_numpy(np.array([1024.0]))
    found = T.zeros(1, "bool")
    prb._C_ops.check_finite_and_unscale_([t], scale, found)
    bits_equal(t.numpy(), x * (1.0 / 1024.0))
    assert not found.numpy()[0]
    x2 = x.copy()
    x2[77] = np.inf
    t2 = T.from_numpy(x2)
    prb._C_ops.check_finite_and_unscale_([t2, t], scale, found)
    assert found.numpy()[0]


def test_amp_scaler_scale_multiplies_by_the_device_loss_scale(prb):
    """AmpScaler.scale(var) = var * loss_scale, a new tensor, no read-back (ref grad_scaler.py:152-200)."""
    scaler = prb.amp.GradScaler(init_loss_scaling=2.0 ** 15)
    loss = prb.Tensor.from_numpy(np.array([0.1234], np.float32))
    out = scaler.scale(loss)
    assert out.data_ptr != loss.data_ptr
    assert out.numpy()[0] == np.float32(0.1234) * np.float32(2.0 ** 15)
    x = orc.from_f32(np.random.default_rng(1).normal(0, 1, 1001).astype(np.float32), orc.BF16)
    got = scaler.scale(prb.Tensor.from_numpy(x, orc.BF16)).numpy()
    bits_equal(got, orc.from_f32(orc.to_f32(x, orc.BF16) * np.float32(2.0 ** 15), orc.BF16))
    assert prb.amp.GradScaler(enable=False).scale(loss) is loss
