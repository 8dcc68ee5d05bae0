"""GPU parity of the CUDA path against three more fixture families generated from the reference's own tests
(tests/golden/make_golden.py):
  * fused_adam_optest.npz      -- test_fused_adam_op.py (inputs exactly as that OpTest builds them) + fused_adam_step
  * adamw_layerwise.npz        -- test_adamw_op.py:614-755, lr_ratio per layer, five steps, rtol 1e-6
  * update_loss_scaling_cases  -- test_update_loss_scaling_op.py:52-136, the two known-answer cases
"""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    return prb


def test_fused_adam_optest_inputs(prb, golden_dir):
    T = prb.Tensor
    npz = np.load(os.path.join(golden_dir, "fused_adam_optest.npz"))
    num = int(npz["num"][0])
    up = lambda k: [T.from_numpy(npz[f"in/{k}/{i}"].ravel().astype(np.float32)) for i in range(num)]  # noqa: E731
    params, grads, m1, m2, b1p, b2p = (up(k) for k in ("Params", "Grads", "Moments1", "Moments2", "Beta1Pows", "Beta2Pows"))
    lr = T.from_numpy(npz["LearningRate"].astype(np.float32))
    beta1, beta2, eps = (float(npz["attr/" + k]) for k in ("beta1", "beta2", "epsilon"))
    prb._C_ops.fused_adam_(params, grads, lr, m1, m2, b1p, b2p, None, None, beta1, beta2, eps,
                           int(npz["attr/chunk_size"]), 0.0, False)
    for i in range(num):
        np.testing.assert_allclose(params[i].numpy(), npz[f"out/ParamsOut/{i}"].ravel(), atol=1e-5, rtol=0)
        np.testing.assert_allclose(m1[i].numpy(), npz[f"out/Moments1Out/{i}"].ravel(), atol=1e-5, rtol=0)
        np.testing.assert_allclose(m2[i].numpy(), npz[f"out/Moments2Out/{i}"].ravel(), atol=1e-5, rtol=0)
        np.testing.assert_allclose(b1p[i].numpy(), npz[f"out/Beta1PowsOut/{i}"].ravel(), rtol=1e-6)
        np.testing.assert_allclose(b2p[i].numpy(), npz[f"out/Beta2PowsOut/{i}"].ravel(), rtol=1e-6)


def test_adamw_layerwise_lr_ratio(prb, golden_dir):
    T = prb.Tensor
    npz = np.load(os.path.join(golden_dir, "adamw_layerwise.npz"))
    lr_v, wd, beta1, beta2, eps = (float(v) for v in npz["hyper"])
    lr = T.from_numpy(np.array([lr_v], np.float32))
    for name in ("linear_1.w_0", "linear_1.b_0", "linear_2.w_0", "linear_2.b_0"):
        ratio = float(npz[f"{name}/lr_ratio"][0])
        p = T.from_numpy(npz[f"{name}/param0"].astype(np.float32).ravel())
        m1, m2 = T.zeros(p.numel, "float32"), T.zeros(p.numel, "float32")
        b1p = T.from_numpy(np.array([beta1], np.float32))
        b2p = T.from_numpy(np.array([beta2], np.float32))
        for t in range(1, int(npz["steps"][0]) + 1):
            g = T.from_numpy(npz[f"{name}/step{t}/Grad"].astype(np.float32).ravel())
            prb._C_ops.adamw_(p, g, lr, m1, m2, b1p, b2p, None, None, beta1, beta2, eps, ratio, wd, True)
            # tolerance: see tests/test_oracle_golden.py::test_adamw_layerwise_lr_ratio_five_steps
            np.testing.assert_allclose(p.numpy(), npz[f"{name}/step{t}/ParamOut"].ravel(), rtol=1e-6, atol=5e-8)


def test_update_loss_scaling_known_answers(prb, golden_dir):
    T = prb.Tensor
    with open(os.path.join(golden_dir, "update_loss_scaling_cases.json")) as f:
        cases = json.load(f)
    for c in cases:
        x = T.from_numpy(np.random.default_rng(3).random(4096).astype(np.float32))
        found = T.from_numpy(np.array([c["found_inf"]], np.bool_))
        scale = T.from_numpy(np.array([c["prev_loss_scaling"]], np.float32))
        good = T.from_numpy(np.array([c["in_good_steps"]], np.int32))
        bad = T.from_numpy(np.array([c["in_bad_steps"]], np.int32))
        prb._C_ops.update_loss_scaling_([x], found, scale, good, bad, c["incr_every_n_steps"],
                                        c["decr_every_n_nan_or_inf"], c["incr_ratio"], c["decr_ratio"])
        assert scale.numpy()[0] == pytest.approx(c["expected_loss_scaling"], rel=1e-6), c["name"]
        assert int(good.numpy()[0]) == c["expected_good_steps"] and int(bad.numpy()[0]) == c["expected_bad_steps"]
        if c["expected_out_zeroed"]:
            assert not x.numpy().any()
