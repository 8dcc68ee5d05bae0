"""The gradient crosses HBM once: (1) the reducer's pack carries check_finite_and_unscale's inf/nan check and the
squared-norm partials (producer-side pass 1), (2) gradients produced directly in the bucket arena need no pack at all.
Both must leave exactly the bits of the three-pass path (pack, norm pass, update pass) and of the oracle."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.test_gpu_tail import SIZES, check, make_params, oracle_tail
from tests.util import bits_equal

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    prb.distributed.init_parallel_env()
    return prb


def _launches(prb):
    return prb._lib.lib.prb_launch_count()


def _setup(prb, seed, dtype, gdt, amp, clip_norm, producer_pass1, arena=False, sizes=SIZES, find_unused=False):
    rng = np.random.default_rng(seed)
    params, host = make_params(prb, rng, sizes, dtype, grad_scale=(2.0 ** 15 * 30 if amp else 30.0), grad_dtype=gdt)
    clip = prb.nn.ClipGradByGlobalNorm(clip_norm) if clip_norm else None
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, weight_decay=0.01, grad_clip=clip,
                              multi_precision=dtype != orc.F32)
    opt._use_producer_pass1 = producer_pass1
    scaler = prb.amp.GradScaler(init_loss_scaling=2.0 ** 15, incr_every_n_steps=3, decr_every_n_nan_or_inf=1) if amp else None
    first = [p.grad.numpy() for p in params]
    if arena:
        for p in params:
            p.grad = None
    dp = prb.DataParallel(params, comm_buffer_size=0.1, last_comm_buffer_size=0.01, consume_from_buckets=True,
                          grads_in_arena=arena, find_unused_parameters=find_unused)
    if arena:
        for p, g in zip(params, first):
            p.grad.copy_from_numpy(g)
    return params, host, clip, opt, scaler, dp, rng


def _run_step(prb, dp, opt, scaler, unused=()):
    dp.prepare_for_backward(unused_vars=unused)
    if unused:
        for i in reversed(range(len(dp._parameters))):
            if i not in unused:
                dp.mark_var_ready(i)
    else:
        dp.mark_all_ready()
    if scaler is not None:
        scaler.step(opt)
        scaler.update()
    else:
        opt.step()


@pytest.mark.parametrize("dtype,gdt,amp,clip", [(orc.BF16, orc.BF16, True, 1.0), (orc.F32, orc.F32, False, 1.0),
                                                (orc.BF16, orc.BF16, True, None), (orc.F16, orc.F16, True, 0.5),
                                                (orc.BF16, orc.F32, True, 1.0)])
def test_pack_epilogue_equals_three_pass_path_and_oracle(prb, dtype, gdt, amp, clip):
    A = _setup(prb, 21, dtype, gdt, amp, clip, producer_pass1=False)
    B = _setup(prb, 21, dtype, gdt, amp, clip, producer_pass1=True)
    pa, ha, ca, oa, sa, da, ra = A
    pb, hb, cb, ob, sb, db, rb = B
    scale, b1p, b2p = 2.0 ** 15, np.float32(0.9), np.float32(0.999)
    fused_launches = plain_launches = None
    for step in range(5):
        if step > 0:
            for (p, q, h) in zip(pa, pb, hb):
                g = orc.from_f32((ra.normal(0, 1e-2, p.numel) * (2.0 ** 15 * 30 if amp else 30.0)).astype(np.float32), gdt)
                if step == 2 and amp and p.name == "w_5":
                    g[17] = orc.from_f32(np.array([np.inf], np.float32), gdt)[0]
                p.grad.copy_from_numpy(g)
                q.grad.copy_from_numpy(g)
                h["grad"] = g
        n0 = _launches(prb)
        _run_step(prb, da, oa, sa)
        n1 = _launches(prb)
        _run_step(prb, db, ob, sb)
        n2 = _launches(prb)
        if step == 0:
            assert not db._epilogue_armed or db._epilogue_key is None  # first step: nothing configured before its backward
        else:
            assert db._epilogue_armed, "the reducer did not carry pass 1"
            fused_launches, plain_launches = n2 - n1, n1 - n0
        for a, b in zip(pa, pb):
            bits_equal(a.numpy(), b.numpy())
            for k in ("moment1", "moment2"):
                bits_equal(oa._accumulators[a.name][k].numpy(), ob._accumulators[b.name][k].numpy())
            if dtype != orc.F32:
                bits_equal(oa._master_weights[a.name].numpy(), ob._master_weights[b.name].numpy())
        if clip:
            bits_equal(ca.last_global_norm.numpy(), cb.last_global_norm.numpy())
            bits_equal(ca.last_clip.numpy(), cb.last_clip.numpy())
        if amp:
            assert sa._found_inf.numpy()[0] == sb._found_inf.numpy()[0] == (step == 2)
            bits_equal(sa._scale.numpy(), sb._scale.numpy())
        # and the oracle's composition of the reference's stages
        found, norm, clipf, _ = oracle_tail(hb, dtype, gdt, loss_scale=scale if amp else None, clip_norm=clip,
                                            b1p=b1p, b2p=b2p,
                                            clip_override=cb.last_clip.numpy()[0] if clip else None)
        assert found == (amp and step == 2)
        if not found:
            check(prb, ob, pb, hb, dtype)
            b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
            if clip:
                tol = 4e-3 if gdt != orc.F32 else 2e-6
                np.testing.assert_allclose(cb.last_global_norm.numpy()[0], norm, rtol=tol)
        elif amp:
            scale = scale / 2
    assert fused_launches == plain_launches - 1, (fused_launches, plain_launches)  # no grad_pass_kernel in the fused step


def test_pack_epilogue_with_unused_parameters_and_reconfiguration(prb):
    """Unused parameters contribute zero to the norm; an explicit scaler.unscale_() turns the epilogue off again."""
    A = _setup(prb, 22, orc.BF16, orc.BF16, True, 1.0, producer_pass1=False, find_unused=True)
    B = _setup(prb, 22, orc.BF16, orc.BF16, True, 1.0, producer_pass1=True, find_unused=True)
    pa, ha, ca, oa, sa, da, ra = A
    pb, hb, cb, ob, sb, db, rb = B
    for step, unused in enumerate([(), (2, 7), (), (5,)]):
        _run_step(prb, da, oa, sa, unused)
        _run_step(prb, db, ob, sb, unused)
        assert db._epilogue_armed == (step > 0)
        bits_equal(ca.last_global_norm.numpy(), cb.last_global_norm.numpy())
        for a, b in zip(pa, pb):
            bits_equal(a.numpy(), b.numpy())
    # explicit unscale_: the check and the norm must see what unscale_ left in the bucket
    for dp, opt, sc in ((da, oa, sa), (db, ob, sb)):
        dp.prepare_for_backward()
        dp.mark_all_ready()
        sc.unscale_(opt)
        sc.step(opt)
        sc.update()
    assert db._epilogue_key is None
    bits_equal(ca.last_global_norm.numpy(), cb.last_global_norm.numpy())
    for a, b in zip(pa, pb):
        bits_equal(a.numpy(), b.numpy())


@pytest.mark.parametrize("dtype,amp,clip", [(orc.F32, False, None), (orc.BF16, True, 1.0)])
def test_gradients_produced_in_the_arena(prb, dtype, amp, clip):
    """grads_in_arena: p.grad IS the bucket slot; a step launches no pack and gives the bits of the packed path
    (up to -0 -> +0 of the scale, which the packed path applies and an untouched slot does not: no -0 in this data)."""
    A = _setup(prb, 23, dtype, dtype, amp, clip, producer_pass1=False)
    B = _setup(prb, 23, dtype, dtype, amp, clip, producer_pass1=True, arena=True)
    pa, ha, ca, oa, sa, da, ra = A
    pb, hb, cb, ob, sb, db, rb = B
    for p, v in zip(pb, db.bucket_grads()):
        assert p.grad.data_ptr == v.data_ptr
    for step in range(3):
        if step > 0:
            for (p, q) in zip(pa, pb):
                g = orc.from_f32((ra.normal(0, 1e-2, p.numel) * (2.0 ** 15 * 30 if amp else 30.0)).astype(np.float32), dtype)
                p.grad.copy_from_numpy(g)
                q.grad.copy_from_numpy(g)
        n0 = _launches(prb)
        _run_step(prb, da, oa, sa)
        n1 = _launches(prb)
        _run_step(prb, db, ob, sb)
        n2 = _launches(prb)
        assert (n2 - n1) == (n1 - n0) - 1, (n2 - n1, n1 - n0)  # the pack launch is gone
        for a, b in zip(pa, pb):
            bits_equal(a.numpy(), b.numpy())
            bits_equal(oa._accumulators[a.name]["moment2"].numpy(), ob._accumulators[b.name]["moment2"].numpy())
        if clip:
            bits_equal(ca.last_clip.numpy(), cb.last_clip.numpy())


@pytest.mark.parametrize("dtype", [orc.BF16, orc.F32])
def test_clip_with_exactly_representable_norm(prb, dtype):
    """Gradients whose squares and partial sums are exact whatever the summation order, so norm and clip factor must
    equal the oracle's BIT FOR BIT.  bf16: two tensors of 4097 ones have per-tensor sums 4097, which the op's T-typed
    output rounds to 4096 (squared_l2_norm_kernel.cu:23-35 writes T); with 2048 twos (8192) the group sum is 16384,
    the norm 128 and the factor 2^-7 -- an implementation that forgot the per-tensor rounding gets sqrt(16386)."""
    sizes = [4097, 4097, 2048]
    vals = [1.0, 1.0, 2.0]
    grads = [orc.from_f32(np.full(n, v, np.float32), dtype) for n, v in zip(sizes, vals)]
    want_norm, want_clip = orc.global_norm_clip([g.copy() for g in grads], [dtype] * 3, 1.0)
    if dtype == orc.BF16:
        assert want_norm == 128.0 and want_clip == 2.0 ** -7
    else:
        assert want_norm == np.float32(np.sqrt(np.float32(16386.0)))
    # (1) the operator
    ts = [prb.Tensor.from_numpy(g, dtype) for g in grads]
    norm, clip = prb._C_ops.clip_grad_by_global_norm_(ts, 1.0)
    assert norm.numpy()[0] == np.float32(want_norm) and clip.numpy()[0] == np.float32(want_clip)
    # (2) the fused step, pass 1 in the tail (step 0) and in the reducer's pack (step 1)
    params = []
    for i, g in enumerate(grads):
        p = prb.Tensor.from_numpy(orc.from_f32(np.zeros(g.size, np.float32), dtype), dtype, name="e_%d" % i)
        p.grad = prb.Tensor.from_numpy(g, dtype)
        params.append(p)
    c = prb.nn.ClipGradByGlobalNorm(1.0)
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, grad_clip=c, multi_precision=dtype != orc.F32)
    dp = prb.DataParallel(params, comm_buffer_size=0.01, last_comm_buffer_size=0.001, consume_from_buckets=True)
    for step in range(2):
        dp.prepare_for_backward()
        dp.mark_all_ready()
        opt.step()
        assert dp._epilogue_armed == (step == 1)
        assert c.last_global_norm.numpy()[0] == np.float32(want_norm), step
        assert c.last_clip.numpy()[0] == np.float32(want_clip), step


@pytest.mark.parametrize("dtype", [orc.BF16, orc.F16])
def test_pack_of_every_16_bit_pattern_equals_the_scale_operator(prb, dtype):
    """The one-rank pack takes a shortcut for 16-bit types (one packed add of +0 instead of unpack / add / multiply /
    repack).  It must give the bits of scale_(x, 1.0, 0.0, False) for EVERY bit pattern -- signed zeros, subnormals,
    infinities, NaNs included -- and the oracle's bits wherever the value is not a NaN."""
    x = np.arange(65536, dtype=np.uint16)
    p = prb.Tensor.from_numpy(x.copy(), dtype, name="allbits")
    p.grad = prb.Tensor.from_numpy(x.copy(), dtype)
    dp = prb.DataParallel([p], consume_from_buckets=True)
    active, _ = dp._reducer.set_grad_epilogue(partial_base=np.zeros(1, np.int32))   # norm partials only
    assert active
    dp.prepare_for_backward()
    dp.mark_all_ready()
    got = dp.bucket_grads()[0].numpy()
    ref = prb._C_ops.scale(p.grad, 1.0, 0.0, False).numpy()        # the element-wise operator kernel
    bits_equal(got, ref)
    want = orc.scale(x, dtype, 1.0)
    finite = np.isfinite(orc.to_f32(x, dtype)) | np.isinf(orc.to_f32(x, dtype))
    bits_equal(got[finite], want[finite])
    assert got[0x8000] == 0                                        # -0 -> +0
