"""GPU parity of the reducer (single rank), the fused step tail and the Python API
(DataParallel / AdamW / GradScaler / ClipGradByGlobalNorm) against the oracle's composition of the
reference's stages in the reference's order."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests.util import bits_equal

pytestmark = pytest.mark.gpu

SIZES = [64, 64, 3 * 3 * 64 * 16, 1000, 7, 4096, 4097, 250000, 1]


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    prb.distributed.init_parallel_env()
    return prb


def make_params(prb, rng, sizes, dtype, grad_scale=1.0, grad_dtype=None):
    grad_dtype = dtype if grad_dtype is None else grad_dtype
    params, host = [], []
    for i, n in enumerate(sizes):
        p = orc.from_f32(rng.normal(0, 0.02, n).astype(np.float32), dtype)
        g = orc.from_f32((rng.normal(0, 1e-2, n) * grad_scale).astype(np.float32), grad_dtype)
        t = prb.Tensor.from_numpy(p, dtype, name="w_%d" % i)
        t.grad = prb.Tensor.from_numpy(g, grad_dtype)
        params.append(t)
        host.append({"param": p.copy(), "grad": g.copy(), "mom1": np.zeros(n, np.float32), "mom2": np.zeros(n, np.float32),
                     "master": orc.to_f32(p, dtype) if dtype != orc.F32 else None})
    return params, host


def oracle_tail(host, dtype, grad_dtype, *, nranks=1, loss_scale=None, clip_norm=None, lr=1e-3, beta1=0.9, beta2=0.999,
                eps=1e-8, coeff=0.01, b1p=0.9, b2p=0.999, adamw=True, lr_ratio=None, with_decay=None,
                clip_override=None):
    """Reference order: concat -> scale(1/nranks) -> [allreduce] -> split -> check_finite_and_unscale ->
    global-norm clip (new tensors) -> adamw per tensor (skipped when found_inf)."""
    grads = [h["grad"] for h in host]
    bucket = orc.scale(orc.concat(grads), grad_dtype, 1.0 / nranks)
    grads = orc.split(bucket, [g.size for g in grads])
    found = False
    if loss_scale is not None:
        found = orc.check_finite_and_unscale(grads, grad_dtype, loss_scale)
    visible = [g.copy() for g in grads]          # what p.grad holds after the step
    norm = clip = None
    if clip_norm is not None:
        # the oracle's own norm / clip factor (compared with tolerance: summation order is unpinned) ...
        norm, clip = orc.global_norm_clip([g.copy() for g in grads], [grad_dtype] * len(grads), clip_norm)
        # ... and the multiply with the factor the device derived, so everything downstream is bit-comparable
        for g in grads:
            orc.clip_mul(g, grad_dtype, float(clip if clip_override is None else clip_override))
    if not found:
        for i, (h, g) in enumerate(zip(host, grads)):
            wd = True if with_decay is None else bool(with_decay[i])
            orc.adam(h["param"], g, h["mom1"], h["mom2"], h["master"], dtype_param=dtype, dtype_grad=grad_dtype, lr=lr,
                     lr_ratio=1.0 if lr_ratio is None else float(lr_ratio[i]), beta1=beta1, beta2=beta2, epsilon=eps,
                     coeff=coeff if (adamw and wd) else 0.0, use_decay=adamw, beta1_pow=np.float32(b1p),
                     beta2_pow=np.float32(b2p))
    return found, norm, clip, visible


def check(prb, opt, params, host, dtype):
    for p, h in zip(params, host):
        bits_equal(p.numpy(), h["param"])
        acc = opt._accumulators[p.name]
        bits_equal(acc["moment1"].numpy(), h["mom1"])
        bits_equal(acc["moment2"].numpy(), h["mom2"])
        if h["master"] is not None:
            bits_equal(opt._master_weights[p.name].numpy(), h["master"])


def test_reducer_single_rank_pack_unpack_is_bit_exact(prb):
    rng = np.random.default_rng(0)
    params, host = make_params(prb, rng, SIZES, orc.F32)
    for p, h in zip(params, host):
        h["grad"][:1] = -0.0
        p.grad.copy_from_numpy(h["grad"])
    dp = prb.DataParallel(params, comm_buffer_size=0.1, last_comm_buffer_size=0.01)
    assert len(dp.group_indices) > 2
    for _ in range(2):
        dp.prepare_for_backward()
        dp.mark_all_ready()
        assert dp.finalized
    for p, h in zip(params, host):
        want = orc.scale(h["grad"], orc.F32, 1.0)  # nranks = 1: only -0 -> +0 changes
        bits_equal(p.grad.numpy(), want)
    # bucket views hold the same reduced values
    dp.consume_from_buckets(True)
    dp.prepare_for_backward()
    dp.mark_all_ready()
    for v, p in zip(dp.bucket_grads(), params):
        bits_equal(v.numpy(), p.grad.numpy())


@pytest.mark.parametrize("consume", [False, True])
def test_dataparallel_adamw_fp32_step(prb, consume):
    """configs[1] shape of the path: DataParallel + AdamW fp32, no AMP, no clip; 3 steps."""
    rng = np.random.default_rng(1)
    params, host = make_params(prb, rng, SIZES, orc.F32)
    dp = prb.DataParallel(params, comm_buffer_size=0.1, last_comm_buffer_size=0.01, consume_from_buckets=consume)
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, weight_decay=0.01,
                              apply_decay_param_fun=lambda name: name not in ("w_0", "w_1"),
                              lr_ratio=lambda p: 0.5 if p.name == "w_3" else 1.0)
    wd = [p.name not in ("w_0", "w_1") for p in params]
    ratio = [0.5 if p.name == "w_3" else 1.0 for p in params]
    b1p, b2p = np.float32(0.9), np.float32(0.999)
    for step in range(3):
        for p, h in zip(params, host):
            g = rng.normal(0, 1e-2, p.numel).astype(np.float32)
            p.grad.copy_from_numpy(g)
            h["grad"] = g
        dp.prepare_for_backward()
        dp.mark_all_ready()
        opt.step()
        oracle_tail(host, orc.F32, orc.F32, b1p=b1p, b2p=b2p, lr_ratio=ratio, with_decay=wd)
        b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
        check(prb, opt, params, host, orc.F32)
    assert opt._beta1_pow.numpy()[0] == b1p and opt._beta2_pow.numpy()[0] == b2p


@pytest.mark.parametrize("main_grad", [False, True])
@pytest.mark.parametrize("inject", [None, "inf", "nan"])
def test_amp_bf16_master_clip_step(prb, main_grad, inject):
    """configs[2] shape of the path: bf16 params + fp32 master, loss scale 2^15, global-norm clip 1.0."""
    rng = np.random.default_rng(2)
    gdt = orc.F32 if main_grad else orc.BF16
    params, host = make_params(prb, rng, SIZES, orc.BF16, grad_scale=2.0 ** 15 * 30, grad_dtype=gdt)
    clip = prb.nn.ClipGradByGlobalNorm(1.0)
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, weight_decay=0.01, grad_clip=clip,
                              multi_precision=True)
    scaler = prb.amp.GradScaler(init_loss_scaling=2.0 ** 15, incr_every_n_steps=2, decr_every_n_nan_or_inf=1)
    dp = prb.DataParallel(params, comm_buffer_size=0.1, last_comm_buffer_size=0.01)
    if inject:
        bad = host[5]["grad"]
        bad[17] = orc.from_f32(np.array([np.inf if inject == "inf" else np.nan], np.float32), gdt)[0]
        params[5].grad.copy_from_numpy(bad)
    dp.prepare_for_backward()
    dp.mark_all_ready()
    scaler.step(opt)
    scaler.update()
    got_clip = clip.last_clip.numpy()[0]
    found, norm, clipf, visible = oracle_tail(host, orc.BF16, gdt, loss_scale=2.0 ** 15, clip_norm=1.0,
                                              clip_override=got_clip)
    assert found == (inject is not None)
    assert bool(scaler._found_inf.numpy()[0]) == found
    if not found:
        tol = 4e-3 if gdt == orc.BF16 else 2e-6   # summation order / T-typed partial sums (SURVEY 8 note 8)
        np.testing.assert_allclose(clip.last_global_norm.numpy()[0], norm, rtol=tol)
        np.testing.assert_allclose(got_clip, clipf, rtol=tol)
        assert got_clip < 1.0  # the test data really is clipped
        check(prb, opt, params, host, orc.BF16)
        for p, v in zip(params, visible):  # p.grad = unscaled, unclipped gradient (reference-visible state)
            bits_equal(p.grad.numpy(), v)
        assert opt._beta1_pow.numpy()[0] == np.float32(0.9) * np.float32(0.9)
    else:
        # skipped step: params, moments, master and beta-pows untouched; loss scale halved
        for p, h in zip(params, host):
            bits_equal(p.numpy(), h["param"])
            bits_equal(opt._master_weights[p.name].numpy(), h["master"])
            assert not opt._accumulators[p.name]["moment1"].numpy().any()
        assert opt._beta1_pow.numpy()[0] == np.float32(0.9)
        assert scaler._scale.numpy()[0] == 2.0 ** 14
    sd = scaler.state_dict()
    assert sd["incr_count"] == (0 if found else 1) and sd["decr_count"] == 0


def test_unscale_then_step_matches_fused_step(prb):
    """scaler.unscale_(opt) followed by scaler.step(opt) (found_inf as an input-only skip flag)."""
    rng = np.random.default_rng(4)
    pa, ha = make_params(prb, rng, SIZES, orc.F32, grad_scale=1024.0)
    rng = np.random.default_rng(4)
    pb, hb = make_params(prb, rng, SIZES, orc.F32, grad_scale=1024.0)
    oa = prb.optimizer.AdamW(learning_rate=1e-3, parameters=pa)
    ob = prb.optimizer.AdamW(learning_rate=1e-3, parameters=pb)
    sa, sb = prb.amp.GradScaler(init_loss_scaling=1024.0), prb.amp.GradScaler(init_loss_scaling=1024.0)
    sa.step(oa)
    sb.unscale_(ob)
    sb.step(ob)
    for a, b in zip(pa, pb):
        bits_equal(a.numpy(), b.numpy())
        bits_equal(a.grad.numpy(), b.grad.numpy())
    oracle_tail(ha, orc.F32, orc.F32, loss_scale=1024.0)
    check(prb, oa, pa, ha, orc.F32)


def test_mixed_dtypes_fall_back_to_the_per_tensor_loop(prb):
    rng = np.random.default_rng(6)
    p32, h32 = make_params(prb, rng, [1000, 33], orc.F32)
    p16, h16 = make_params(prb, rng, [4097], orc.BF16)
    p16[0].name = "w_bf16"
    params = p32 + p16
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, multi_precision=True)
    opt.step()
    oracle_tail(h32, orc.F32, orc.F32)
    oracle_tail(h16, orc.BF16, orc.BF16)
    check(prb, opt, params, h32 + h16, None)
    assert opt._beta1_pow.numpy()[0] == np.float32(0.9) * np.float32(0.9)


def test_adam_use_multi_tensor_and_state_dict_roundtrip(prb):
    rng = np.random.default_rng(9)
    params, host = make_params(prb, rng, [500, 4096, 77], orc.F32)
    opt = prb.optimizer.Adam(learning_rate=1e-3, parameters=params, use_multi_tensor=True)
    opt.step()
    oracle_tail(host, orc.F32, orc.F32, adamw=False)
    check(prb, opt, params, host, orc.F32)
    sd = opt.state_dict()
    assert set(sd) == {f"w_{i}_{k}_0" for i in range(3) for k in ("moment1", "moment2", "beta1_pow_acc", "beta2_pow_acc")}
    opt2 = prb.optimizer.Adam(learning_rate=1e-3, parameters=params)
    opt2.set_state_dict(sd)
    for p in params:
        bits_equal(opt2._accumulators[p.name]["moment1"].numpy(), opt._accumulators[p.name]["moment1"].numpy())
    assert opt2._beta1_pow.numpy()[0] == opt._beta1_pow.numpy()[0]
    opt.clear_grad()
    assert not params[1].grad.numpy().any()


def test_clear_grad_folded_into_the_step(prb):
    rng = np.random.default_rng(10)
    params, host = make_params(prb, rng, SIZES, orc.F32)
    opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params)
    opt._zero_grads_in_step = True
    opt.step()
    oracle_tail(host, orc.F32, orc.F32)
    check(prb, opt, params, host, orc.F32)
    for p in params:
        assert not p.grad.numpy().any()


def test_fleet_style_fused_allreduce_single_rank(prb):
    """hybrid_parallel_util.py:73-111 with one rank: scale = 1.0 -> no scale, grads come back bit for bit;
    an explicit scale goes through Tensor.scale_ (scale * x + 0); main_grad takes precedence over grad."""
    rng = np.random.default_rng(12)
    params, host = make_params(prb, rng, SIZES, orc.BF16)
    params[2].main_grad = prb.Tensor.from_numpy(rng.normal(0, 1, params[2].numel).astype(np.float32))
    mg = params[2].main_grad.numpy()
    params[2].grad = None
    prb.distributed.fused_allreduce_gradients_with_group(params, None, bucket_size=1 << 20)
    for i, (p, h) in enumerate(zip(params, host)):
        if i == 2:
            bits_equal(p.main_grad.numpy(), mg)
        else:
            bits_equal(p.grad.numpy(), h["grad"])
    prb.distributed.fused_allreduce_gradients_with_group(params, None, bucket_size=1 << 20, scale=0.25)
    for i, (p, h) in enumerate(zip(params, host)):
        if i == 2:
            bits_equal(p.main_grad.numpy(), orc.scale(mg, orc.F32, 0.25, 0.0, True))
        else:
            bits_equal(p.grad.numpy(), orc.scale(h["grad"], orc.BF16, 0.25, 0.0, True))


def test_reducer_hooks_in_any_order_unused_params_and_no_sync(prb):
    """EagerReducer on the device: hooks arriving out of order (groups still launch in index order,
    reducer.cc:916-919), parameters without a gradient this step (zero-filled slots, :833-855) and no_sync."""
    rng = np.random.default_rng(13)
    params, host = make_params(prb, rng, SIZES, orc.F32)
    dp = prb.DataParallel(params, comm_buffer_size=0.1, last_comm_buffer_size=0.01, find_unused_parameters=True)
    ngroups = len(dp.group_indices)
    assert ngroups >= 3
    # 1) random hook order
    dp.prepare_for_backward()
    order = list(rng.permutation(len(params)))
    for k, i in enumerate(order):
        dp.mark_var_ready(int(i))
        assert dp._reducer.launched_groups() == sorted(dp._reducer.launched_groups())
        assert dp.finalized == (k == len(order) - 1)
    assert dp._reducer.launched_groups() == list(range(ngroups))
    for p, h in zip(params, host):
        bits_equal(p.grad.numpy(), orc.scale(h["grad"], orc.F32, 1.0))
    # 2) two parameters get no gradient this step: their slots are zero-filled, everything else is untouched
    unused = [2, 7]
    dp.prepare_for_backward(unused_vars=unused)
    for i in reversed(range(len(params))):
        if i not in unused:
            dp.mark_var_ready(i)
    assert dp.finalized
    for i, (p, h) in enumerate(zip(params, host)):
        if i in unused:
            assert not p.grad.numpy().any()
            p.grad.copy_from_numpy(h["grad"])
        else:
            bits_equal(p.grad.numpy(), orc.scale(h["grad"], orc.F32, 1.0))
    # 3) no_sync: hooks do nothing, the next synchronised backward works as usual
    with dp.no_sync():
        dp.prepare_for_backward()
        dp.mark_all_ready()
    dp.prepare_for_backward()
    dp.mark_all_ready()
    assert dp.finalized
    for p, h in zip(params, host):
        bits_equal(p.grad.numpy(), orc.scale(h["grad"], orc.F32, 1.0))


def test_group_sharded_optimizer_on_one_rank_matches_the_oracle(prb):
    """GroupShardedOptimizerStage2 with a single rank owns everything: the tail split in two calls (check + norm +
    clip factor over all gradients, then the update with the parameter written over its gradient) must give the
    bits of the one-call tail -- the AMP bf16 + master + clip job of bench.parity_preamble, gradients packed and
    produced in the arena.  (Several ranks: tests/dist_worker.py.)"""
    import bench
    group = prb.distributed.init_parallel_env()
    for arena in (False, True):
        for layout in (True, "bucket"):
            assert bench.parity_preamble(prb, group, arena=arena, sharded=layout) == "ok"
