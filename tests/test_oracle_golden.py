"""Pins the CPU oracle against the reference's own test oracles / literal vectors.

Fixtures come from tests/golden/make_golden.py, which executed the reference's numpy
helpers (adamw_step, adam_step) and read the literal bucket cases from the
reference's tests.  Tolerances are the reference's: OpTest.check_output default
atol=1e-5 (eager_op_test.py:2036-2038); multi-step adam rtol 1e-6 is *not* claimed by
the reference for adam_step (other factoring, SURVEY 8 parity note 4), so adam uses
the OpTest default too.
"""
import json
import os

import numpy as np
import pytest

from oracle import oracle as orc


def _case(npz, prefix):
    ins = {k.split("/")[-1]: npz[k] for k in npz.files if k.startswith(prefix + "/in/")}
    attrs = {k.split("/")[-1]: npz[k].item() for k in npz.files if k.startswith(prefix + "/attr/")}
    outs = {k.split("/")[-1]: npz[k] for k in npz.files if k.startswith(prefix + "/out/")}
    return ins, attrs, outs


def _run_oracle_adam(ins, attrs, adamw, fma=False):
    p = ins["Param"].astype(np.float32).ravel().copy()
    g = ins["Grad"].astype(np.float32).ravel().copy()
    m1 = ins["Moment1"].astype(np.float32).ravel().copy()
    m2 = ins["Moment2"].astype(np.float32).ravel().copy()
    orc.adam(p, g, m1, m2, None, dtype_param=orc.F32, dtype_grad=orc.F32, lr=float(ins["LearningRate"][0]),
             lr_ratio=float(attrs.get("lr_ratio", 1.0)), beta1=attrs["beta1"], beta2=attrs["beta2"],
             epsilon=attrs["epsilon"], coeff=float(attrs.get("coeff", 0.0)) if attrs.get("with_decay", False) else 0.0,
             use_decay=adamw, beta1_pow=float(ins["Beta1Pow"][0]), beta2_pow=float(ins["Beta2Pow"][0]), fma=fma)
    return p, m1, m2


@pytest.mark.parametrize("name", ["TestAdamW", "TestAdamW2"])
@pytest.mark.parametrize("fma", [False, True])
def test_adamw_matches_reference_numpy_oracle(golden_dir, name, fma):
    npz = np.load(os.path.join(golden_dir, "adamw_optest.npz"))
    ins, attrs, outs = _case(npz, name)
    p, m1, m2 = _run_oracle_adam(ins, attrs, adamw=True, fma=fma)
    np.testing.assert_allclose(p, outs["ParamOut"].ravel(), atol=1e-5, rtol=0)
    np.testing.assert_allclose(m1, outs["Moment1Out"].ravel(), atol=1e-5, rtol=0)
    np.testing.assert_allclose(m2, outs["Moment2Out"].ravel(), atol=1e-5, rtol=0)
    # adamw_step has the kernel's algebra (SURVEY 8 note 4): relative agreement is much tighter
    np.testing.assert_allclose(p, outs["ParamOut"].ravel(), rtol=2e-6, atol=1e-7)


@pytest.mark.parametrize("name", ["TestAdamOp1", "TestAdamOp2", "TestAdamOnlyTailOp"])
def test_adam_matches_reference_numpy_oracle(golden_dir, name):
    npz = np.load(os.path.join(golden_dir, "adam_optest.npz"))
    ins, attrs, outs = _case(npz, name)
    p, m1, m2 = _run_oracle_adam(ins, attrs, adamw=False)
    np.testing.assert_allclose(p, outs["ParamOut"].ravel(), atol=1e-5, rtol=0)
    np.testing.assert_allclose(m1, outs["Moment1Out"].ravel(), atol=1e-5, rtol=0)
    np.testing.assert_allclose(m2, outs["Moment2Out"].ravel(), atol=1e-5, rtol=0)


def test_adam_multiple_steps(golden_dir):
    npz = np.load(os.path.join(golden_dir, "adam_optest.npz"))
    ins, attrs, _ = _case(npz, "MultipleSteps")
    p = ins["Param"].ravel().copy()
    m1 = ins["Moment1"].ravel().copy()
    m2 = ins["Moment2"].ravel().copy()
    b1p = ins["Beta1Pow"].copy()
    b2p = ins["Beta2Pow"].copy()
    for step in range(10):
        g = npz[f"MultipleSteps/step{step}/Grad"].ravel().copy()
        orc.adam(p, g, m1, m2, None, dtype_param=orc.F32, dtype_grad=orc.F32, lr=float(ins["LearningRate"][0]),
                 beta1=attrs["beta1"], beta2=attrs["beta2"], epsilon=attrs["epsilon"], use_decay=False,
                 beta1_pow=float(b1p[0]), beta2_pow=float(b2p[0]))
        orc.update_beta_pow(attrs["beta1"], attrs["beta2"], b1p, b2p)
        np.testing.assert_allclose(b1p, npz[f"MultipleSteps/step{step}/Beta1PowOut"], rtol=1e-6)
        np.testing.assert_allclose(b2p, npz[f"MultipleSteps/step{step}/Beta2PowOut"], rtol=1e-6)
        if step in (0, 9):
            np.testing.assert_allclose(p, npz[f"MultipleSteps/step{step}/ParamOut"].ravel(), atol=1e-5, rtol=0)
            np.testing.assert_allclose(m1, npz[f"MultipleSteps/step{step}/Moment1Out"].ravel(), atol=1e-5, rtol=0)
            np.testing.assert_allclose(m2, npz[f"MultipleSteps/step{step}/Moment2Out"].ravel(), atol=1e-5, rtol=0)


def test_fma_flavour_within_north_star_tolerance():
    """Contracted vs strictly-rounded evaluation differ by < 1e-6 relative on params."""
    rng = np.random.default_rng(7)
    n = 1 << 16
    p0 = rng.normal(0, 0.02, n).astype(np.float32)
    g = rng.normal(0, 1e-2, n).astype(np.float32)
    m1_0 = rng.normal(0, 1e-3, n).astype(np.float32)
    m2_0 = rng.uniform(0, 1e-3, n).astype(np.float32)
    res = []
    for fma in (False, True):
        p, m1, m2 = p0.copy(), m1_0.copy(), m2_0.copy()
        orc.adam(p, g, m1, m2, None, dtype_param=orc.F32, dtype_grad=orc.F32, lr=1e-3, beta1=0.9, beta2=0.999,
                 epsilon=1e-8, coeff=0.01, use_decay=True, beta1_pow=0.9 ** 10, beta2_pow=0.999 ** 10, fma=fma)
        res.append((p, m1, m2))
    np.testing.assert_allclose(res[0][0], res[1][0], rtol=1e-6, atol=1e-9)
    # moments: beta*m + (1-beta)*g can cancel, so the bound is relative to the operands' scale (~1e-3)
    np.testing.assert_allclose(res[0][1], res[1][1], rtol=1e-6, atol=1e-6 * 1e-3)
    np.testing.assert_allclose(res[0][2], res[1][2], rtol=1e-6, atol=1e-6 * 1e-3)


def test_check_finite_and_unscale_semantics(golden_dir):
    fx = np.load(os.path.join(golden_dir, "check_finite_optest.npz"))
    np.random.seed(123)  # regenerate the reference's input from its seed
    x = np.random.random((1024, 1024)).astype(np.float32)
    scale = np.random.random((1)).astype(np.float32)
    assert scale[0] == fx["scale"][0]
    np.testing.assert_array_equal(x[::16, ::16], fx["x_sample"])
    for kind, expect in (("clean", 0), ("nan", 1), ("inf", 1)):
        xi = x.copy()
        if kind != "clean":
            i, j = fx["poison_index"]
            xi[i, j] = np.nan if kind == "nan" else np.inf
        flat = xi.ravel()
        found = orc.check_finite_and_unscale([flat], orc.F32, float(scale[0]))
        assert int(found) == expect == int(fx[f"found_{kind}"][0])
        if kind == "clean":
            # reference expectation: Out = x / scale at OpTest tolerance; kernel multiplies by
            # the reciprocal computed in double (amp_kernel.cu:29), so allow 1 ulp relative too.
            np.testing.assert_allclose(flat.reshape(1024, 1024)[::16, ::16], fx["out_sample"], rtol=2e-7, atol=1e-5)


def test_check_finite_empty_and_ragged():
    a = np.array([], dtype=np.float32)
    b = np.array([1.0, 2.0, 3.0], dtype=np.float32)
    c = np.array([], dtype=np.float32)
    d = np.array([4.0], dtype=np.float32)
    assert not orc.check_finite_and_unscale([a, b, c, d], orc.F32, 2.0)
    np.testing.assert_array_equal(b, [0.5, 1.0, 1.5])
    np.testing.assert_array_equal(d, [2.0])


def test_check_finite_is_on_narrowed_value():
    # 3e38 * (1/0.5) overflows fp32 -> inf after the multiply; bf16 near-max rounds to inf when narrowed
    x = np.array([3.0e38], dtype=np.float32)
    assert orc.check_finite_and_unscale([x], orc.F32, 0.5)
    big = np.array([0x7F7F], dtype=np.uint16)  # bf16 max = 3.3895e38
    assert np.isfinite(orc.to_f32(big, orc.BF16)[0])
    xb = big.copy()
    # 3.3895e38 / 0.997 = 3.3997e38: finite in fp32, but rounds to +inf when narrowed to bf16
    assert np.isfinite(np.float32(orc.to_f32(big, orc.BF16)[0]) * np.float32(1.0 / 0.997))
    assert orc.check_finite_and_unscale([xb], orc.BF16, 0.997)
    assert xb[0] == 0x7F80


def test_squared_l2_norm(golden_dir):
    fx = np.load(os.path.join(golden_dir, "squared_l2_norm_optest.npz"))
    X = fx["TestL2LossOp/X"].ravel().copy()
    v, v64 = orc.squared_l2_norm(X, orc.F32)
    np.testing.assert_allclose(v, fx["TestL2LossOp/Out"][0], rtol=1e-5)
    np.testing.assert_allclose(v64, fx["TestL2LossOp/Out"][0], rtol=1e-5)
    for name in ("F16Op", "F16Op1", "F16Op2"):
        x32 = fx[f"{name}/X"].ravel().copy()
        x16 = orc.from_f32(x32, orc.F16)
        v32, _ = orc.squared_l2_norm(x32, orc.F32)
        # reference compares fp16 run with fp32 run at rtol 1e-5 -- which only holds with
        # fp32 inputs cast on the fly; with the T-rounded squares the agreement is ~1e-3.
        v16, _ = orc.squared_l2_norm(x16, orc.F16)
        np.testing.assert_allclose(v16, v32, rtol=2e-3)
    # determinism (test_squared_l2_norm_op.py:98-111)
    assert orc.squared_l2_norm(X, orc.F32) == orc.squared_l2_norm(X, orc.F32)


def test_bucket_assignment_golden(golden_dir):
    names = {"float32": orc.F32, "float64": orc.F64, "float16": orc.F16, "bfloat16": orc.BF16}
    cases = json.load(open(os.path.join(golden_dir, "bucket_cases.json")))
    assert len(cases) == 10
    for c in cases:
        numels = [int(np.prod(s)) for s in c["shapes"]]
        dts = [names[d] for d in c["dtypes"]]
        got = orc.assign_group_by_size(numels, dts, c["is_sparse"], c["limits"], c["tensor_indices"])
        assert got == c["expected"], c["name"]


def test_scale_semantics():
    x = np.array([-0.0, 1.0, -3.0, 1e-45], dtype=np.float32)
    out = orc.scale(x, orc.F32, 1.0 / 1)
    assert np.signbit(out[0]) == False  # noqa: E712  (-0 + 0 -> +0, SURVEY parity note 1)
    out8 = orc.scale(x, orc.F32, 1.0 / 8)
    np.testing.assert_array_equal(out8[1:3], np.array([0.125, -0.375], dtype=np.float32))
    out3 = orc.scale(np.array([1.0], dtype=np.float32), orc.F32, 1.0 / 3)
    assert out3[0] == np.float32(1.0 / 3)


def test_concat_split_roundtrip_bit_exact():
    rng = np.random.default_rng(0)
    sizes = [0, 1, 7, 1024, 3, 0, 65537]
    arrs = [rng.integers(0, 65536, n).astype(np.uint16) for n in sizes]
    bucket = orc.concat(arrs)
    assert bucket.size == sum(sizes)
    outs = orc.split(bucket, sizes)
    for a, o in zip(arrs, outs):
        np.testing.assert_array_equal(a, o)


def test_bf16_fp16_conversions_against_numpy():
    rng = np.random.default_rng(1)
    x = np.concatenate([
        rng.normal(0, 1, 4096).astype(np.float32),
        np.array([0.0, -0.0, np.inf, -np.inf, 65504.0, 65520.0, 65519.9, 6e-8, 5.96e-8, 2.98e-8, 2.99e-8, 1e-10, 6.1e-5],
                 dtype=np.float32),
    ])
    h = orc.from_f32(x, orc.F16)
    np.testing.assert_array_equal(h, x.astype(np.float16).view(np.uint16))
    np.testing.assert_array_equal(orc.to_f32(h, orc.F16), x.astype(np.float16).astype(np.float32))
    b = orc.from_f32(x, orc.BF16)
    # RNE by integer arithmetic on the bit pattern
    u = x.view(np.uint32).astype(np.uint64)
    expect = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    np.testing.assert_array_equal(b, expect)
    assert orc.from_f32(np.array([np.nan], dtype=np.float32), orc.BF16)[0] == 0x7FFF


def test_reducer_model_orders_groups():
    groups = [[5, 4], [3, 2], [1, 0]]
    m = orc.ReducerModel(groups, 6)
    # group 1 completes first but must wait for group 0 (reducer.cc:916-919)
    for v in (3, 2):
        m.mark_var_ready(v)
    assert m.launched == []
    m.mark_var_ready(5)
    m.mark_var_ready(4)
    assert m.launched == [0, 1]
    m.mark_var_ready(1)
    with pytest.raises(RuntimeError):
        m.mark_var_ready(1)
    m.mark_var_ready(0)
    assert m.launched == [0, 1, 2] and m.finalized
    m.prepare_for_backward()
    m.mark_var_ready(0)
    with pytest.raises(RuntimeError):
        m.prepare_for_backward()


def test_update_loss_scaling_state_machine():
    ls, good, bad = 32768.0, 0, 0
    for _ in range(999):
        ls, good, bad = orc.update_loss_scaling(False, ls, good, bad, 1000, 2, 2.0, 0.5)
    assert (ls, good, bad) == (32768.0, 999, 0)
    ls, good, bad = orc.update_loss_scaling(False, ls, good, bad, 1000, 2, 2.0, 0.5)
    assert (ls, good, bad) == (65536.0, 0, 0)
    ls, good, bad = orc.update_loss_scaling(True, ls, good, bad, 1000, 2, 2.0, 0.5)
    assert (ls, good, bad) == (65536.0, 0, 1)
    ls, good, bad = orc.update_loss_scaling(True, ls, good, bad, 1000, 2, 2.0, 0.5)
    assert (ls, good, bad) == (32768.0, 0, 0)
    ls, good, bad = orc.update_loss_scaling(True, 1.5, 0, 1, 1000, 2, 2.0, 0.5)
    assert ls == 1.0


# ------------------------------------------------------------------ further fixtures from the reference's tests
def test_fused_adam_matches_reference_numpy_oracle(golden_dir):
    """test_fused_adam_op.py: the operator's OpTest inputs (as that test really builds them, aliasing included)
    and its numpy oracle; OpTest tolerance."""
    npz = np.load(os.path.join(golden_dir, "fused_adam_optest.npz"))
    num = int(npz["num"][0])
    attrs = {k: float(npz["attr/" + k]) for k in ("epsilon", "beta1", "beta2")}
    lr = float(npz["LearningRate"][0])
    for i in range(num):
        ins = {"Param": npz[f"in/Params/{i}"], "Grad": npz[f"in/Grads/{i}"], "Moment1": npz[f"in/Moments1/{i}"],
               "Moment2": npz[f"in/Moments2/{i}"], "LearningRate": np.array([lr], np.float32),
               "Beta1Pow": npz[f"in/Beta1Pows/{i}"], "Beta2Pow": npz[f"in/Beta2Pows/{i}"]}
        p, m1, m2 = _run_oracle_adam(ins, attrs, adamw=False)
        np.testing.assert_allclose(p, npz[f"out/ParamsOut/{i}"].ravel(), atol=1e-5, rtol=0)
        np.testing.assert_allclose(m1, npz[f"out/Moments1Out/{i}"].ravel(), atol=1e-5, rtol=0)
        np.testing.assert_allclose(m2, npz[f"out/Moments2Out/{i}"].ravel(), atol=1e-5, rtol=0)
        # beta-pow outputs: in * beta, rounded to fp32 (ref fused_adam_kernel.cu:230-238)
        b1 = np.float32(ins["Beta1Pow"][0]) * np.float32(attrs["beta1"])
        np.testing.assert_allclose(b1, npz[f"out/Beta1PowsOut/{i}"][0], rtol=1e-6)


def test_adamw_layerwise_lr_ratio_five_steps(golden_dir):
    """test_adamw_op.py:628-755: lr_ratio per layer, five steps, the reference's own tolerance (rtol 1e-6)."""
    npz = np.load(os.path.join(golden_dir, "adamw_layerwise.npz"))
    lr, wd, beta1, beta2, eps = (float(v) for v in npz["hyper"])
    for name in ("linear_1.w_0", "linear_1.b_0", "linear_2.w_0", "linear_2.b_0"):
        ratio = float(npz[f"{name}/lr_ratio"][0])
        p = npz[f"{name}/param0"].astype(np.float32).ravel().copy()
        m1 = np.zeros_like(p)
        m2 = np.zeros_like(p)
        for t in range(1, int(npz["steps"][0]) + 1):
            g = npz[f"{name}/step{t}/Grad"].astype(np.float32).ravel().copy()
            orc.adam(p, g, m1, m2, None, dtype_param=orc.F32, dtype_grad=orc.F32, lr=float(np.float32(lr)),
                     lr_ratio=ratio, beta1=beta1, beta2=beta2, epsilon=eps, coeff=wd, use_decay=True,
                     beta1_pow=float(np.float32(beta1 ** t)), beta2_pow=float(np.float32(beta2 ** t)))
            # rtol is the reference's.  atol: adamw_step multiplies by float32(1 - beta2) = 0.001, the kernel (and
            # the oracle) by 1.0f - float32(beta2) = 0.00100004673 (adamw_kernel.cu:65-66); that moves every update
            # by ~2e-5 of its size (5e-4 here), which only shows on the few weights that happen to be ~1e-4 small
            np.testing.assert_allclose(p, npz[f"{name}/step{t}/ParamOut"].ravel(), rtol=1e-6, atol=5e-8)
    # the depth rule itself (test_adamw_op.py:614-622): decay_rate ** (n_layers + 2 - depth)
    assert float(npz["linear_1.w_0/lr_ratio"][0]) == pytest.approx(0.8 ** 3)
    assert float(npz["linear_2.b_0/lr_ratio"][0]) == pytest.approx(0.8 ** 2)


def test_update_loss_scaling_known_answers(golden_dir):
    """test_update_loss_scaling_op.py:52-136: the two OpTest cases, literal expectations."""
    import json
    with open(os.path.join(golden_dir, "update_loss_scaling_cases.json")) as f:
        cases = json.load(f)
    assert [c["name"] for c in cases] == ["TestUpdateLossScalingOp", "TestUpdateLossScalingOpBad"]
    for c in cases:
        ls, good, bad = orc.update_loss_scaling(c["found_inf"], c["prev_loss_scaling"], c["in_good_steps"],
                                                c["in_bad_steps"], c["incr_every_n_steps"],
                                                c["decr_every_n_nan_or_inf"], c["incr_ratio"], c["decr_ratio"])
        assert ls == pytest.approx(c["expected_loss_scaling"], rel=1e-6), c["name"]
        assert (good, bad) == (c["expected_good_steps"], c["expected_bad_steps"]), c["name"]


def test_global_norm_clip_property_of_the_reference_test():
    """test_gradient_clip.py:442-482: ||clip(g)|| == min(||g||, clip_norm) (rtol 1e-6, atol 1e-8) on a Linear(5, 5)."""
    rng = np.random.default_rng(7)
    for scale in (10.0, 0.01):  # one case that clips, one that does not
        grads = [(rng.normal(0, scale, 25)).astype(np.float32), (rng.normal(0, scale, 5)).astype(np.float32)]
        before = np.sqrt(sum(np.sum(np.power(g.astype(np.float64), 2)) for g in grads))
        out = [g.copy() for g in grads]
        orc.global_norm_clip(out, [orc.F32, orc.F32], 0.8)  # in place
        after = np.sqrt(sum(np.sum(np.power(np.asarray(g, np.float64), 2)) for g in out))
        assert np.isclose(min(before, 0.8), after, rtol=1e-6, atol=1e-8), (before, after)


# ------------------------------------------------------------------ the timed CPU baseline computes the right thing
def test_cpu_baseline_port_agrees_with_the_gpu_semantics():
    """The CPU path (cpu/adamw_kernel.cc:96-189, refer.h:539-563) factors the update differently from the GPU
    kernel but is the same mathematics: the baseline that bench.py times must land within a few ulp
    of the parity oracle, and must not depend on the thread count (ref jit/test.cc:690-900 holds every CPU
    implementation to the refer one the same way)."""
    rng = np.random.default_rng(5)
    n = 512 * 37 + 129  # whole chunks plus a tail (cpu/adamw_kernel.cc:147-189)
    p0 = rng.normal(0, 0.02, n).astype(np.float32)
    g = rng.normal(0, 1e-2, n).astype(np.float32)
    m10 = rng.normal(0, 1e-3, n).astype(np.float32)
    m20 = rng.uniform(0, 1e-3, n).astype(np.float32)
    kw = dict(lr=1e-3, lr_ratio=1.0, beta1=0.9, beta2=0.999, epsilon=1e-8, coeff=0.01,
              beta1_pow=float(np.float32(0.9 ** 10)), beta2_pow=float(np.float32(0.999 ** 10)))
    outs = []
    for threads in (1, max(2, orc.max_threads())):
        p, m1, m2 = p0.copy(), m10.copy(), m20.copy()
        orc.cpu_adamw(p, g, m1, m2, nthreads=threads, **kw)
        outs.append((p, m1, m2))
    for a, b in zip(*outs):
        np.testing.assert_array_equal(a, b)  # element-wise work: the thread count cannot matter
    p, m1, m2 = p0.copy(), m10.copy(), m20.copy()
    orc.adam(p, g.copy(), m1, m2, None, dtype_param=orc.F32, dtype_grad=orc.F32, use_decay=True, **kw)
    np.testing.assert_allclose(outs[0][0], p, rtol=2e-6, atol=1e-9)
    np.testing.assert_allclose(outs[0][1], m1, rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(outs[0][2], m2, rtol=1e-6, atol=1e-12)
    x = rng.normal(0, 1, 1000).astype(np.float32)
    x[0] = -0.0
    y = x.copy()
    orc.cpu_scale_f32(y, np.float32(0.25), 2)
    np.testing.assert_array_equal(y, np.float32(0.25) * (x + np.float32(0)))


@pytest.mark.parametrize("dt", [orc.BF16, orc.F16])
@pytest.mark.parametrize("main_grad", [False, True])
def test_multi_precision_master_follows_the_fp32_run(dt, main_grad):
    """test_adamw_op.py:323-459: with fp32 master weights the master after the step equals an fp32 AdamW on the
    same numbers (rtol 1e-6 there; exactly here), and the 16-bit parameter is its rounding (rtol 1e-2 there)."""
    rng = np.random.default_rng(11)
    n = 1000
    master0 = rng.normal(0, 1, n).astype(np.float32)
    g32 = rng.normal(0, 1, n).astype(np.float32)
    g16 = orc.from_f32(g32, dt)
    g_as_f32 = g32 if main_grad else orc.to_f32(g16, dt)
    m1 = rng.normal(0, 1, n).astype(np.float32)
    m2 = np.abs(rng.normal(0, 1, n)).astype(np.float32)
    kw = dict(lr=1e-8, lr_ratio=1.0, beta1=0.9, beta2=0.99, epsilon=1e-8, coeff=0.1, use_decay=True,
              beta1_pow=float(np.float32(0.9 ** 10)), beta2_pow=float(np.float32(0.99 ** 10)))
    ref_p, ref_m1, ref_m2 = master0.copy(), m1.copy(), m2.copy()
    orc.adam(ref_p, g_as_f32.copy(), ref_m1, ref_m2, None, dtype_param=orc.F32, dtype_grad=orc.F32, **kw)
    p16 = orc.from_f32(master0, dt)
    master, a1, a2 = master0.copy(), m1.copy(), m2.copy()
    grad = g32.copy() if main_grad else g16.copy()
    orc.adam(p16, grad, a1, a2, master, dtype_param=dt, dtype_grad=orc.F32 if main_grad else dt, **kw)
    np.testing.assert_array_equal(master, ref_p)
    np.testing.assert_array_equal(a1, ref_m1)
    np.testing.assert_array_equal(a2, ref_m2)
    np.testing.assert_array_equal(p16, orc.from_f32(master, dt))
    np.testing.assert_allclose(orc.to_f32(p16, dt), ref_p, rtol=1e-2)


def test_sparse_adam_against_the_reference_optest(golden_dir):
    """TestSparseAdamOp (test_adam_op.py:355-460): adam with a SelectedRows gradient, both lazy modes, against the
    reference's own adam_step_sparse on its literal inputs, with that test's own acceptance criterion."""
    z = np.load(os.path.join(golden_dir, "sparse_adam_optest.npz"))
    beta1, beta2, eps = (float(v) for v in z["hyper"])
    rows = z["rows"]
    for tag, lazy in (("full", False), ("lazy", True)):
        p, m1, m2 = (z["in/" + k].copy() for k in ("Param", "Moment1", "Moment2"))
        orc.sparse_adam(p, rows, z["grad"], m1, m2, None, dtype_param=orc.F32, lr=float(z["in/LearningRate"][0]),
                        beta1=beta1, beta2=beta2, epsilon=eps, beta1_pow=z["in/Beta1Pow"][0],
                        beta2_pow=z["in/Beta2Pow"][0], lazy_mode=lazy)
        touched = rows if lazy else np.arange(p.shape[0])
        for got, key in ((p, "/ParamOut"), (m1, "/Moment1Out"), (m2, "/Moment2Out")):
            want = z[tag + key][touched]
            # the reference's own criterion (test_adam_op.py:442-443) is one-sided: actual - expected < 1e-5
            assert ((got[touched] - want) < 1e-5).all(), key
            # and two-sided: its numpy oracle scales epsilon differently from its kernels (SURVEY 8 note 4); with this
            # test's epsilon = 1e-4 and lr = 2.0 the two factorings are 1.7e-5 apart on the parameters
            np.testing.assert_allclose(got[touched], want, atol=2e-5, rtol=0)
        if lazy:
            untouched = np.setdiff1d(np.arange(p.shape[0]), rows)
            assert (p[untouched] == 5.0).all() and (m1[untouched] == 5.0).all()
    # duplicates are merged before the update (MergeAdd)
    ur, mv = orc.merge_rows([4, 0, 4], np.array([[1.0, 2.0], [3.0, 4.0], [10.0, 20.0]], np.float32), orc.F32)
    assert ur.tolist() == [0, 4] and mv.tolist() == [[3.0, 4.0], [11.0, 22.0]]


def test_sparse_allreduce_and_empty_rows():
    """AllReduceSparse (reducer.cc:1072-1277): values scaled by 1/nranks, rows and values concatenated in rank order,
    whether or not the ranks hold the same number of rows."""
    a = (np.array([3, 1]), np.array([[2.0, 4.0], [6.0, -0.0]], np.float32))
    b = (np.array([1, 7, 7]), np.array([[1.0, 1.0], [2.0, 2.0], [8.0, 8.0]], np.float32))
    rows, vals = orc.sparse_allreduce([a, b], orc.F32)
    assert rows.tolist() == [3, 1, 1, 7, 7]
    assert vals.tolist() == [[1.0, 2.0], [3.0, 0.0], [0.5, 0.5], [1.0, 1.0], [4.0, 4.0]]
    ur, mv = orc.merge_rows(rows, vals, orc.F32)
    assert ur.tolist() == [1, 3, 7] and mv.tolist() == [[3.5, 0.5], [1.0, 2.0], [5.0, 5.0]]
    # no rows at all: the full update still runs with g = 0 everywhere, the lazy one touches nothing
    p = np.full((4, 3), 1.0, np.float32)
    m1, m2 = np.full((4, 3), 0.5, np.float32), np.full((4, 3), 0.25, np.float32)
    kw = dict(dtype_param=orc.F32, lr=0.1, beta1=0.9, beta2=0.999, epsilon=1e-8, beta1_pow=np.float32(0.9),
              beta2_pow=np.float32(0.999))
    q, a1, a2 = p.copy(), m1.copy(), m2.copy()
    orc.sparse_adam(q, np.array([], np.int64), np.zeros((0, 3), np.float32), a1, a2, None, lazy_mode=True, **kw)
    assert (q == p).all() and (a1 == m1).all()
    orc.sparse_adam(q, np.array([], np.int64), np.zeros((0, 3), np.float32), a1, a2, None, lazy_mode=False, **kw)
    assert (a1 == np.float32(0.9) * m1).all() and (q != p).all()
