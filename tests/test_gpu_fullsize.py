"""Full-size runs (BASELINE.json configs 3-5) checked through size-independent properties.

The oracle cannot finish these sizes in seconds, so: tensors are filled with one value per tensor ON THE DEVICE,
the kernels run at full size, and the result must (a) equal the oracle's single-element result bit for bit at the
head, the middle and the very last element of every tensor (catches indexing / 64-bit offset / tail bugs), and
(b) have the sum of squares n * v^2 (every element was touched exactly once).  concat -> split must round-trip.
"""
import numpy as np
import pytest

import bench_configs
from oracle import oracle as orc
from tests.util import bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    return prb


def fill(prb, t, value):
    """t[:] = value, on the device: zero, then scale with bias (0 * 1 + value)."""
    if t.dtype == orc.F32:
        t.zero_()
        prb._C_ops.scale_(t, 1.0, float(value), True)
    else:
        tmp = prb.Tensor.zeros(t.numel, "float32")
        prb._C_ops.scale_(tmp, 1.0, float(value), True)
        t.copy_(tmp.astype(t.dtype))
    return t


def probe(t, n):
    """head, middle and last element of a tensor as a small host array."""
    idx = sorted({0, n // 2, n - 1})
    return np.concatenate([t.view(i, 1).numpy() for i in idx])


def oracle_one(dtype, gdtype, p, g, m1, m2, *, master, lr, coeff, b1p, b2p):
    st_p = orc.from_f32(np.array([p], np.float32), dtype)
    st_g = orc.from_f32(np.array([g], np.float32), gdtype)
    a, b = np.array([m1], np.float32), np.array([m2], np.float32)
    mw = orc.to_f32(st_p, dtype) if master else None
    orc.adam(st_p, st_g, a, b, mw, dtype_param=dtype, dtype_grad=gdtype, lr=lr, beta1=0.9, beta2=0.999, epsilon=1e-8,
             coeff=coeff, use_decay=True, beta1_pow=np.float32(b1p), beta2_pow=np.float32(b2p))
    return st_p[0], a[0], b[0], (mw[0] if master else None)


def test_gpt2_medium_fused_adam_full_size(prb):
    """configs[3]: GPT-2 medium shapes (292 tensors, 354,823,168 elements), bf16 params + fp32 master."""
    numels = bench_configs.param_numels("gpt2_medium")
    T = prb.Tensor
    params, grads, m1s, m2s, masters, vals = [], [], [], [], [], []
    rng = np.random.default_rng(3)
    for i, n in enumerate(numels):
        pv, gv, av, bv = rng.normal(0, 0.02), rng.normal(0, 1e-2), rng.normal(0, 1e-3), rng.uniform(1e-6, 1e-3)
        p = fill(prb, T(n, "bfloat16"), pv)
        params.append(p)
        masters.append(p.astype("float32"))
        grads.append(fill(prb, T(n, "bfloat16"), gv))
        m1s.append(fill(prb, T(n, "float32"), av))
        m2s.append(fill(prb, T(n, "float32"), bv))
        vals.append((pv, gv, av, bv))
    lr = T.from_numpy(np.array([1e-3], np.float32))
    b1 = [T.from_numpy(np.array([0.9 ** 10], np.float32)) for _ in numels]
    b2 = [T.from_numpy(np.array([0.999 ** 10], np.float32)) for _ in numels]
    prb._C_ops.fused_adam_(params, grads, lr, m1s, m2s, b1, b2, masters, None, 0.9, 0.999, 1e-8, 4096, 0.01, True, True, False)
    for i, n in enumerate(numels):
        pv, gv, av, bv = vals[i]
        # the oracle starts from the same rounded-to-bf16 parameter / gradient the device holds
        wp, wa, wb, wm = oracle_one(orc.BF16, orc.BF16, pv, gv, np.float32(av), np.float32(bv), master=True, lr=1e-3,
                                    coeff=0.01, b1p=np.float32(0.9 ** 10), b2p=np.float32(0.999 ** 10))
        assert (probe(params[i], n) == wp).all(), i
        assert (probe(m1s[i], n) == wa).all() and (probe(m2s[i], n) == wb).all(), i
        assert (probe(masters[i], n) == wm).all(), i
        if i % 37 == 0 or n > 10 ** 7:  # every element touched once: sum of squares = n * v^2
            got = prb._C_ops.squared_l2_norm(masters[i]).numpy()[0]
            np.testing.assert_allclose(got, n * float(wm) ** 2, rtol=2e-5)
    assert b1[5].numpy()[0] == np.float32(0.9) * np.float32(0.9 ** 10)


def test_bucket_roundtrip_full_size(prb):
    """configs[3]: 710 MB of bf16 gradients through concat -> split comes back bit for bit."""
    numels = bench_configs.param_numels("gpt2_medium")
    rng = np.random.default_rng(4)
    T = prb.Tensor
    srcs = []
    for n in numels:
        srcs.append(T.from_numpy(rng.integers(0, 65536, n, dtype=np.uint16), "bfloat16"))
    bucket = prb._C_ops.concat(srcs)
    outs = [T.zeros(n, "bfloat16") for n in numels]
    prb._C_ops.split_(bucket, outs)
    for i in (0, 1, 2, 57, 150, 291):
        bits_equal(outs[i].numpy(), srcs[i].numpy())
    whole = bucket.numpy()
    off = 0
    for i, n in enumerate(numels):
        if i in (0, 2, 291):
            bits_equal(whole[off:off + n], srcs[i].numpy())
        off += n


def test_two_billion_element_sweep_fp32(prb):
    """configs[4] upper end: one fp32 tensor of 2*10^9 elements (the reference indexes with int, adamw_kernel.cu:50,55)."""
    n = 2_000_000_000
    T = prb.Tensor
    pv, gv, av, bv = 0.0123, -0.0071, 3.0e-4, 5.0e-4
    p, g, m1, m2 = (fill(prb, T(n, "float32"), v) for v in (pv, gv, av, bv))
    lr = T.from_numpy(np.array([1e-3], np.float32))
    b1, b2 = T.from_numpy(np.array([0.9 ** 10], np.float32)), T.from_numpy(np.array([0.999 ** 10], np.float32))
    # the very last element lives past 2^31 elements and past 2^33 bytes
    assert (n - 1) > 2 ** 30 and (n - 1) * 4 > 2 ** 32
    found = T.zeros(1, "bool")
    scale = T.from_numpy(np.array([4.0], np.float32))
    prb._C_ops.check_finite_and_unscale_([g], scale, found)
    assert not found.numpy()[0] and (probe(g, n) == np.float32(gv) * np.float32(0.25)).all()
    norm = prb._C_ops.squared_l2_norm(g).numpy()[0]
    np.testing.assert_allclose(norm, n * float(np.float32(gv) * np.float32(0.25)) ** 2, rtol=2e-5)
    prb._C_ops.adamw_(p, g, lr, m1, m2, b1, b2, None, None, 0.9, 0.999, 1e-8, 1.0, 0.01, True)
    wp, wa, wb, _ = oracle_one(orc.F32, orc.F32, np.float32(pv), np.float32(gv) * np.float32(0.25), np.float32(av),
                               np.float32(bv), master=False, lr=1e-3, coeff=0.01, b1p=np.float32(0.9 ** 10),
                               b2p=np.float32(0.999 ** 10))
    assert (probe(p, n) == wp).all() and (probe(m1, n) == wa).all() and (probe(m2, n) == wb).all()
    np.testing.assert_allclose(prb._C_ops.squared_l2_norm(p).numpy()[0], n * float(wp) ** 2, rtol=2e-5)
    # a poisoned LAST element is found
    g.view(n - 1, 1).copy_from_numpy(np.array([np.inf], np.float32))
    prb._C_ops.check_finite_and_unscale_([g], scale, found)
    assert found.numpy()[0]


def _full_config_step(prb, shapes_name, dtype, amp, clip_norm, steps):
    """The bench workload at its real size through the public API, every tensor checked against the oracle's
    composition of the reference's stages (concat -> scale -> split -> unscale/check -> clip -> adamw)."""
    numels = bench_configs.param_numels(shapes_name)
    rng = np.random.default_rng(77)
    params, host = [], []
    for i, n in enumerate(numels):
        p = orc.from_f32(rng.standard_normal(n, dtype=np.float32) * np.float32(0.02), dtype)
        t = prb.Tensor.from_numpy(p, dtype, name="w_%d" % i)
        t.grad = prb.Tensor.zeros(n, dtype)
        params.append(t)
        host.append({"param": p, "mom1": np.zeros(n, np.float32), "mom2": np.zeros(n, np.float32),
                     "master": orc.to_f32(p, dtype) if dtype != orc.F32 else None})
    clip = prb.nn.ClipGradByGlobalNorm(clip_norm) if clip_norm else None
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, weight_decay=0.01, grad_clip=clip,
                              multi_precision=dtype != orc.F32)
    scaler = prb.amp.GradScaler(init_loss_scaling=2.0 ** 15) if amp else None
    dp = prb.DataParallel(params, comm_buffer_size=25, last_comm_buffer_size=1, consume_from_buckets=True)
    b1p, b2p = np.float32(0.9), np.float32(0.999)
    for step in range(steps):
        grads = []
        for p in params:
            g = rng.standard_normal(p.numel, dtype=np.float32) * np.float32(1e-2 * (2.0 ** 15 if amp else 1.0))
            g = orc.from_f32(g, dtype)
            p.grad.copy_from_numpy(g)
            grads.append(g)
        dp.prepare_for_backward()
        dp.mark_all_ready()
        if scaler is not None:
            scaler.step(opt)
            scaler.update()
        else:
            opt.step()
        # ---- oracle, reference order
        bucket = orc.scale(orc.concat(grads), dtype, 1.0)
        grads = orc.split(bucket, numels)
        if amp:
            assert not orc.check_finite_and_unscale(grads, dtype, 2.0 ** 15)
            assert not scaler._found_inf.numpy()[0]
        if clip_norm:
            norm, cl = orc.global_norm_clip([g.copy() for g in grads], [dtype] * len(grads), clip_norm)
            got_clip = clip.last_clip.numpy()[0]
            tol = 4e-3 if dtype != orc.F32 else 2e-6
            np.testing.assert_allclose(clip.last_global_norm.numpy()[0], norm, rtol=tol)
            np.testing.assert_allclose(got_clip, cl, rtol=tol)
            for g in grads:
                orc.clip_mul(g, dtype, float(got_clip))
        for h, g in zip(host, grads):
            orc.adam(h["param"], g, h["mom1"], h["mom2"], h["master"], dtype_param=dtype, dtype_grad=dtype, lr=1e-3,
                     beta1=0.9, beta2=0.999, epsilon=1e-8, coeff=0.01, use_decay=True, beta1_pow=b1p, beta2_pow=b2p)
        b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
    for p, h in zip(params, host):
        bits_equal(p.numpy(), h["param"])
        acc = opt._accumulators[p.name]
        bits_equal(acc["moment1"].numpy(), h["mom1"])
        bits_equal(acc["moment2"].numpy(), h["mom2"])
        if h["master"] is not None:
            bits_equal(opt._master_weights[p.name].numpy(), h["master"])
    assert opt._beta1_pow.numpy()[0] == b1p


def test_resnet50_config_bit_exact_vs_oracle(prb):
    """configs[1] at full size: 161 tensors, 25,557,032 fp32 elements, DataParallel + AdamW, 2 steps."""
    prb.distributed.init_parallel_env()
    _full_config_step(prb, "resnet50", orc.F32, amp=False, clip_norm=None, steps=2)


def test_bert_base_config_bit_exact_vs_oracle(prb):
    """configs[2] at full size: 206 tensors, 110,106,428 bf16 elements + fp32 master, AMP (2^15) + clip 1.0."""
    prb.distributed.init_parallel_env()
    _full_config_step(prb, "bert_base", orc.BF16, amp=True, clip_norm=1.0, steps=1)
