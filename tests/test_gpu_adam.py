"""GPU parity of adamw_ / adam_ / merged_adam_ / fused_adam_ through the C ABI (via the ctypes shim).

Bars: fp32 update <= 1e-6 relative vs the oracle -- the kernels evaluate the same
individually-rounded expression tree, so the tests assert BIT equality; bf16 outputs are the RNE
rounding of the fp32-master result (0 ulp here, <= 1 ulp required).
"""
import os

import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import bits_equal, copy_state, make_state

pytestmark = pytest.mark.gpu

HP = dict(lr=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8)
B1P, B2P = np.float32(0.9 ** 10), np.float32(0.999 ** 10)


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    return prb


def upload(prb, st, dtype_param, dtype_grad):
    T = prb.Tensor
    d = {
        "param": T.from_numpy(st["param"], dtype_param),
        "grad": T.from_numpy(st["grad"], dtype_grad),
        "mom1": T.from_numpy(st["mom1"]),
        "mom2": T.from_numpy(st["mom2"]),
        "master": None if st["master"] is None else T.from_numpy(st["master"]),
    }
    return d


def scalars(prb, lr=HP["lr"], b1p=B1P, b2p=B2P, host_pows=False):
    T = prb.Tensor
    lr_t = T.from_numpy(np.array([lr], np.float32))
    dev = -1 if host_pows else None
    return lr_t, T.from_numpy(np.array([b1p], np.float32), device=dev), T.from_numpy(np.array([b2p], np.float32), device=dev)


def oracle_step(st, dtype_param, dtype_grad, *, use_decay, coeff=0.0, lr_ratio=1.0, lr=HP["lr"], b1p=B1P, b2p=B2P,
                beta1=HP["beta1"], beta2=HP["beta2"], epsilon=HP["epsilon"]):
    orc.adam(st["param"], st["grad"], st["mom1"], st["mom2"], st["master"], dtype_param=dtype_param,
             dtype_grad=dtype_grad, lr=lr, lr_ratio=lr_ratio, beta1=beta1, beta2=beta2, epsilon=epsilon, coeff=coeff,
             use_decay=use_decay, beta1_pow=b1p, beta2_pow=b2p)


def check_state(dev, st):
    bits_equal(dev["param"].numpy(), st["param"])
    bits_equal(dev["mom1"].numpy(), st["mom1"])
    bits_equal(dev["mom2"].numpy(), st["mom2"])
    if st["master"] is not None:
        bits_equal(dev["master"].numpy(), st["master"])


@pytest.mark.parametrize("n", [1, 7, 8, 9, 4095, 4096, 4097, 102 * 105, (1 << 20) + 3])
@pytest.mark.parametrize("with_decay", [True, False])
def test_adamw_fp32_bit_exact(prb, n, with_decay):
    rng = np.random.default_rng(1234 + n)
    st = make_state(rng, n)
    dev = upload(prb, st, orc.F32, orc.F32)
    lr, b1p, b2p = scalars(prb)
    prb._C_ops.adamw_(dev["param"], dev["grad"], lr, dev["mom1"], dev["mom2"], b1p, b2p, None, None, HP["beta1"],
                      HP["beta2"], HP["epsilon"], 0.5, 0.01, with_decay)
    oracle_step(st, orc.F32, orc.F32, use_decay=True, coeff=0.01 if with_decay else 0.0, lr_ratio=0.5)
    check_state(dev, st)
    # device-resident beta-pows advance on the device (adamw_kernel.cu:332-341)
    assert b1p.numpy()[0] == np.float32(0.9) * B1P and b2p.numpy()[0] == np.float32(0.999) * B2P


@pytest.mark.parametrize("n", [3, 102 * 105, 70001])
def test_adam_fp32_bit_exact_host_and_device_beta_pows(prb, n):
    rng = np.random.default_rng(n)
    st = make_state(rng, n)
    for host_pows in (True, False):
        s = copy_state(st)
        dev = upload(prb, s, orc.F32, orc.F32)
        lr, b1p, b2p = scalars(prb, host_pows=host_pows)
        prb._C_ops.adam_(dev["param"], dev["grad"], lr, dev["mom1"], dev["mom2"], b1p, b2p, None, None, HP["beta1"],
                         HP["beta2"], HP["epsilon"])
        oracle_step(s, orc.F32, orc.F32, use_decay=False)
        check_state(dev, s)
        assert b1p.numpy()[0] == np.float32(0.9) * B1P  # host accumulators advance on the host (:283-289)


@pytest.mark.parametrize("dtype", [orc.BF16, orc.F16])
@pytest.mark.parametrize("main_grad", [False, True])
@pytest.mark.parametrize("n", [5, 4096 + 13, 300000])
def test_adamw_multi_precision_16bit(prb, dtype, main_grad, n):
    """16-bit param + fp32 master (+ optional fp32 main_grad): test_adamw_op.py:323-459."""
    rng = np.random.default_rng(n + dtype)
    dtg = orc.F32 if main_grad else dtype
    st = make_state(rng, n, dtype_param=dtype, dtype_grad=dtg, master=True)
    dev = upload(prb, st, dtype, dtg)
    lr, b1p, b2p = scalars(prb)
    prb._C_ops.adamw_(dev["param"], dev["grad"], lr, dev["mom1"], dev["mom2"], b1p, b2p, dev["master"], None,
                      HP["beta1"], HP["beta2"], HP["epsilon"], 1.0, 0.1, True, False, 1000, True, False)
    oracle_step(st, dtype, dtg, use_decay=True, coeff=0.1)
    check_state(dev, st)
    # the 16-bit output is the RNE rounding of the fp32-master result: 0 ulp (<= 1 ulp required)
    bits_equal(dev["param"].numpy(), orc.from_f32(dev["master"].numpy(), dtype))


def test_adamw_bf16_without_master(prb):
    rng = np.random.default_rng(5)
    st = make_state(rng, 12345, dtype_param=orc.BF16)
    dev = upload(prb, st, orc.BF16, orc.BF16)
    lr, b1p, b2p = scalars(prb)
    prb._C_ops.adamw_(dev["param"], dev["grad"], lr, dev["mom1"], dev["mom2"], b1p, b2p, None, None, HP["beta1"],
                      HP["beta2"], HP["epsilon"], 1.0, 0.01, True)
    oracle_step(st, orc.BF16, orc.BF16, use_decay=True, coeff=0.01)
    check_state(dev, st)


def test_unaligned_pointers_take_the_scalar_path(prb):
    """A view starting 1 element into an allocation is only element-aligned."""
    rng = np.random.default_rng(9)
    n = 10001
    st = make_state(rng, n)
    T = prb.Tensor
    big = {k: T.zeros(n + 8, orc.F32) for k in ("param", "grad", "mom1", "mom2")}
    dev = {k: big[k].view(1, n) for k in big}
    for k in big:
        dev[k].copy_from_numpy(st[k])
    dev["master"] = None
    lr, b1p, b2p = scalars(prb)
    prb._C_ops.adamw_(dev["param"], dev["grad"], lr, dev["mom1"], dev["mom2"], b1p, b2p, None, None, HP["beta1"],
                      HP["beta2"], HP["epsilon"], 1.0, 0.01, True)
    oracle_step(st, orc.F32, orc.F32, use_decay=True, coeff=0.01)
    check_state(dev, st)
    for k in big:  # neighbours untouched
        full = big[k].numpy()
        assert full[0] == 0 and np.all(full[n + 1:] == 0)


def test_skip_update_on_device_is_a_no_op(prb):
    rng = np.random.default_rng(11)
    st = make_state(rng, 5000)
    dev = upload(prb, st, orc.F32, orc.F32)
    lr, b1p, b2p = scalars(prb)
    skip = prb.Tensor.from_numpy(np.array([True]))
    prb._C_ops.adamw_(dev["param"], dev["grad"], lr, dev["mom1"], dev["mom2"], b1p, b2p, None, skip, HP["beta1"],
                      HP["beta2"], HP["epsilon"], 1.0, 0.01, True)
    check_state(dev, st)
    assert b1p.numpy()[0] == B1P and b2p.numpy()[0] == B2P
    skip.copy_from_numpy(np.array([False]))
    prb._C_ops.adamw_(dev["param"], dev["grad"], lr, dev["mom1"], dev["mom2"], b1p, b2p, None, skip, HP["beta1"],
                      HP["beta2"], HP["epsilon"], 1.0, 0.01, True)
    oracle_step(st, orc.F32, orc.F32, use_decay=True, coeff=0.01)
    check_state(dev, st)


def test_merged_adam_equals_per_tensor_adam_bitwise(prb):
    """test_merged_adam_op.py:204-217 (assert_array_equal on GPU), plus the oracle."""
    rng = np.random.default_rng(10)
    sizes = [2 * 2, 102 * 105, 1, 33 * 7]
    T = prb.Tensor
    states = [make_state(rng, n) for n in sizes]
    merged = [upload(prb, s, orc.F32, orc.F32) for s in states]
    single = [upload(prb, s, orc.F32, orc.F32) for s in states]
    lrs = [T.from_numpy(np.array([1e-3 * (i + 1)], np.float32)) for i in range(len(sizes))]
    mk = lambda v: [T.from_numpy(np.array([v], np.float32)) for _ in sizes]  # noqa: E731
    b1m, b2m, b1s, b2s = mk(B1P), mk(B2P), mk(B1P), mk(B2P)
    prb._C_ops.merged_adam_([d["param"] for d in merged], [d["grad"] for d in merged], lrs,
                            [d["mom1"] for d in merged], [d["mom2"] for d in merged], b1m, b2m, None, 0.9, 0.999, 1e-8)
    for i, d in enumerate(single):
        prb._C_ops.adam_(d["param"], d["grad"], lrs[i], d["mom1"], d["mom2"], b1s[i], b2s[i], None, None, 0.9, 0.999, 1e-8)
    for i, (m, s) in enumerate(zip(merged, single)):
        for k in ("param", "mom1", "mom2"):
            bits_equal(m[k].numpy(), s[k].numpy())
        oracle_step(states[i], orc.F32, orc.F32, use_decay=False, lr=np.float32(1e-3 * (i + 1)))
        check_state(m, states[i])
        assert b1m[i].numpy()[0] == b1s[i].numpy()[0] == np.float32(0.9) * B1P


@pytest.mark.parametrize("use_adamw", [False, True])
@pytest.mark.parametrize("mp", [False, True])
def test_fused_adam_equals_per_tensor_over_5_steps(prb, use_adamw, mp):
    """test/cpp/phi/kernels/test_fused_adam_kernel.cc:391-468: 40 tensors of random length 0..524288
    (zero-length allowed), 5 steps, beta=(.9,.99), wd .1, eps 1e-6, lr 1e-3, atol 0."""
    rng = np.random.default_rng(10)
    ntens = 40 if os.environ.get("PRB_FULL", "0") == "1" else 12
    sizes = [int(x) for x in rng.integers(0, 524288, ntens)]
    sizes[3] = 0
    dt = orc.BF16 if mp else orc.F32
    T = prb.Tensor
    states = [make_state(rng, n, dtype_param=dt, master=mp) for n in sizes]
    fused = [upload(prb, s, dt, dt) for s in states]
    single = [upload(prb, s, dt, dt) for s in states]
    lr = T.from_numpy(np.array([1e-3], np.float32))
    mk = lambda v: [T.from_numpy(np.array([v], np.float32)) for _ in sizes]  # noqa: E731
    b1f, b2f, b1s, b2s = mk(0.9), mk(0.99), mk(0.9), mk(0.99)
    b1p, b2p = np.float32(0.9), np.float32(0.99)
    for step in range(5):
        grads = [orc.from_f32(rng.normal(0, 1, n).astype(np.float32), dt) for n in sizes]
        for i in range(len(sizes)):
            fused[i]["grad"].copy_from_numpy(grads[i])
            single[i]["grad"].copy_from_numpy(grads[i])
            states[i]["grad"] = grads[i]
        prb._C_ops.fused_adam_([d["param"] for d in fused], [d["grad"] for d in fused], lr, [d["mom1"] for d in fused],
                               [d["mom2"] for d in fused], b1f, b2f, [d["master"] for d in fused] if mp else None, None,
                               0.9, 0.99, 1e-6, 4096, 0.1, use_adamw, mp, False)
        for i, d in enumerate(single):
            if use_adamw:
                prb._C_ops.adamw_(d["param"], d["grad"], lr, d["mom1"], d["mom2"], b1s[i], b2s[i], d["master"], None,
                                  0.9, 0.99, 1e-6, 1.0, 0.1, True, False, 1000, mp, False)
            else:
                prb._C_ops.adam_(d["param"], d["grad"], lr, d["mom1"], d["mom2"], b1s[i], b2s[i], d["master"], None,
                                 0.9, 0.99, 1e-6, False, 1000, mp, False)
            if sizes[i]:
                oracle_step(states[i], dt, dt, use_decay=use_adamw, coeff=0.1, b1p=b1p, b2p=b2p, beta2=0.99, epsilon=1e-6)
        b1p, b2p = np.float32(0.9) * b1p, np.float32(0.99) * b2p
    for i in range(len(sizes)):
        for k in ("param", "mom1", "mom2") + (("master",) if mp else ()):
            bits_equal(fused[i][k].numpy(), single[i][k].numpy())  # atol = 0
        if sizes[i]:
            check_state(fused[i], states[i])
        assert b1f[i].numpy()[0] == b1s[i].numpy()[0] == b1p


def test_golden_fixtures_from_reference_tests(prb, golden_dir):
    """The reference's own OpTest cases (inputs from its seed, expectations from its numpy oracle)."""
    T = prb.Tensor
    for fname, names, adamw in (("adamw_optest.npz", ["TestAdamW", "TestAdamW2"], True),
                                ("adam_optest.npz", ["TestAdamOp1", "TestAdamOp2", "TestAdamOnlyTailOp"], False)):
        npz = np.load(os.path.join(golden_dir, fname))
        for name in names:
            g = lambda k: npz[f"{name}/{k}"]  # noqa: E731
            attr = lambda k, d=None: npz[f"{name}/attr/{k}"].item() if f"{name}/attr/{k}" in npz.files else d  # noqa: E731
            p, gr = T.from_numpy(g("in/Param").ravel()), T.from_numpy(g("in/Grad").ravel())
            m1, m2 = T.from_numpy(g("in/Moment1").ravel()), T.from_numpy(g("in/Moment2").ravel())
            lr, b1p, b2p = (T.from_numpy(g(f"in/{k}")) for k in ("LearningRate", "Beta1Pow", "Beta2Pow"))
            if adamw:
                prb._C_ops.adamw_(p, gr, lr, m1, m2, b1p, b2p, None, None, attr("beta1"), attr("beta2"), attr("epsilon"),
                                  attr("lr_ratio", 1.0), attr("coeff"), attr("with_decay"))
            else:
                prb._C_ops.adam_(p, gr, lr, m1, m2, b1p, b2p, None, None, attr("beta1"), attr("beta2"), attr("epsilon"))
            np.testing.assert_allclose(p.numpy(), g("out/ParamOut").ravel(), atol=1e-5, rtol=0)  # OpTest default
            np.testing.assert_allclose(m1.numpy(), g("out/Moment1Out").ravel(), atol=1e-5, rtol=0)
            np.testing.assert_allclose(m2.numpy(), g("out/Moment2Out").ravel(), atol=1e-5, rtol=0)
            np.testing.assert_allclose(b1p.numpy(), g("in/Beta1Pow") * np.float32(attr("beta1")), rtol=1e-6)


def test_error_behaviour_mirrors_enforce_messages(prb):
    T = prb.Tensor
    from paddle_ray_b200._lib import PRBError
    p, g, m1, m2 = (T.zeros(16, orc.F32) for _ in range(4))
    lr = T.from_numpy(np.array([1e-3], np.float32))
    bad_pow = T.zeros(2, orc.F32)
    ok_pow = T.zeros(1, orc.F32)
    with pytest.raises(PRBError, match="beta1 pow output size should be 1"):
        prb._C_ops.adamw_(p, g, lr, m1, m2, bad_pow, ok_pow, None, None, 0.9, 0.999, 1e-8, 1.0, 0.01, True)
    with pytest.raises(PRBError, match="SkipUpdate"):
        prb._C_ops.adamw_(p, g, lr, m1, m2, ok_pow, ok_pow, None, T.zeros(2, "bool"), 0.9, 0.999, 1e-8, 1.0, 0.01, True)
    host_pow = T.zeros(1, orc.F32, device=-1)
    with pytest.raises(PRBError, match="same place"):
        prb._C_ops.adamw_(p, g, lr, m1, m2, host_pow, ok_pow, None, None, 0.9, 0.999, 1e-8, 1.0, 0.01, True)
