"""``paddle.optimizer.Adam`` / ``AdamW`` host glue over the multi-tensor kernels.

ref: python/paddle/optimizer/optimizer.py:294-409 (state_dict), :414-474 (global lr tensor),
     :640-666 (_create_master_weight), :849-1036 (_create_optimization_pass), :1178-1221
     (_apply_optimize), :1333-1381 (clear_grad), :1447-1504 (step);
     python/paddle/optimizer/adam.py:163-224,530-712; python/paddle/optimizer/adamw.py:149-470.

B200-first differences (results unchanged):
  * moments / master weights live in flat arenas (one allocation per kind), per-parameter views
    are 256-byte aligned so every access is a 256-bit vector access;
  * beta-pow accumulators live on the DEVICE and are shared by all parameters of the optimizer (the
    reference keeps one CPU pair per parameter, all holding the same value); ``state_dict`` still
    emits the reference's per-parameter keys;
  * ``found_inf`` is consumed on the device: no D2H sync, no host branch (ref optimizer.py:983-989);
  * with uniform dtypes, ``step()`` is ONE ``prb_dp_step_tail`` call (unscale + inf check + global
    norm clip + update, two passes) instead of 3P+P+... launches.
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict

import numpy as np

from .. import _C_ops, _lib, core
from .._lib import BFLOAT16, BOOL, FLOAT16, FLOAT32, TailConfig, call, lib, tensor_array
from ..core import SelectedRows, Tensor
from ..nn.clip import ClipGradByGlobalNorm, _clip_by_global_norm_using_mp_type
from .lr import LRScheduler

_ALIGN_ELEMS = 64  # 256 B of fp32
_MISSING = object()  # a cached plan may legitimately be None (the per-tensor path)


class _ScalarFeeder:
    """Asynchronous host->device updates of one float (the learning rate) without a stream sync."""

    SLOTS = 64

    def __init__(self):
        self.buf = core.PinnedBuffer(self.SLOTS, np.float32)
        self.i = 0

    def feed(self, tensor: Tensor, value: float, stream=None):
        self.buf.array[self.i] = np.float32(value)
        call("prb_memcpy_h2d", tensor.data_ptr, self.buf.ptr + 4 * self.i, 4, core._stream_handle(stream))
        self.i = (self.i + 1) % self.SLOTS


class Optimizer:
    _moment1_acc_str = "moment1"
    _moment2_acc_str = "moment2"
    _beta1_pow_acc_str = "beta1_pow_acc"
    _beta2_pow_acc_str = "beta2_pow_acc"

    def __init__(self, learning_rate, parameters, weight_decay=None, grad_clip=None, name=None,
                 multi_precision=False):
        if parameters is None:
            raise AttributeError("parameters argument given to the Optimizer should not be None in dygraph mode.")
        if isinstance(parameters, (Tensor, dict)):
            raise TypeError("`parameters` argument given to the optimizer should be an iterable of Tensors")
        if not isinstance(learning_rate, (float, LRScheduler)):
            raise TypeError("learning rate should be float or LRScheduler, got %s here" % type(learning_rate))
        if grad_clip is not None and not isinstance(grad_clip, ClipGradByGlobalNorm):
            raise TypeError("'grad_clip' should be an instance of ClipGradByGlobalNorm (the clip on this path)")
        self._parameter_list = list(parameters)
        if self._parameter_list and isinstance(self._parameter_list[0], dict):
            raise _lib.PRBError(4, "parameter groups are not built on this path; pass a flat parameter list")
        for i, p in enumerate(self._parameter_list):
            if p.name is None:
                p.name = "param_%d" % i
        self._learning_rate = learning_rate
        self._grad_clip = grad_clip
        self._multi_precision = multi_precision
        self._name = name
        self._master_weights = {}
        self._accumulators = {}
        self._auxiliary_vars = {}
        self._created = False
        self._lr_tensor = None
        self._lr_value = None
        self._lr_feeder = None
        self._step_plans = {}
        self._launch_events = None      # pool of (start, stop) Event pairs handed in by a benchmark
        self._launch_events_used = []
        # True folds clear_grad(set_to_zero=True) into the update pass (gradients are zeroed as they are consumed)
        self._zero_grads_in_step = False
        # True keeps p.grad reference-visible after an AMP step (the unscaled gradient is stored back)
        self._write_back_unscaled_grads = True
        # True lets the producer of the bucket (DataParallel's pack / exchange kernel) carry the inf check and the
        # norm partials, so the step reads each gradient once; False keeps pass 1 here (used by parity tests)
        self._use_producer_pass1 = True
        self.clear_gradients = self.clear_grad

    # ---- auxiliary vars (found_inf hand-off from GradScaler; ref adamw.py:283-291)
    def _set_auxiliary_var(self, key, val):
        self._auxiliary_vars[key] = val

    def _get_auxiliary_var(self, key):
        return self._auxiliary_vars.get(key)

    # ---- learning rate (ref optimizer.py:414-474,505-585)
    def get_lr(self) -> float:
        return self._learning_rate() if isinstance(self._learning_rate, LRScheduler) else float(self._learning_rate)

    def set_lr(self, value: float):
        if isinstance(self._learning_rate, LRScheduler):
            raise RuntimeError("optimizer's learning rate can't be LRScheduler when invoke this API")
        self._learning_rate = float(value)

    def _global_lr(self) -> Tensor:
        lr = self.get_lr()
        if self._lr_tensor is None:
            self._lr_tensor = Tensor.from_numpy(np.array([lr], np.float32))
            self._lr_feeder = _ScalarFeeder()
            self._lr_value = lr
        elif lr != self._lr_value:
            self._lr_feeder.feed(self._lr_tensor, lr)
            self._lr_value = lr
        return self._lr_tensor

    # ---- accumulators
    @staticmethod
    def _is_dtype_fp16_or_bf16(dtype):
        return dtype in (FLOAT16, BFLOAT16)

    def _arena(self, params, dtype):
        offs, total = [], 0
        for p in params:
            offs.append(total)
            total += (p.numel + _ALIGN_ELEMS - 1) // _ALIGN_ELEMS * _ALIGN_ELEMS
        arena = Tensor.zeros(max(total, 1), dtype)
        return arena, [arena.view(o, p.numel) for o, p in zip(offs, params)]

    def _create_accumulators(self):
        if self._created:
            return
        ps = self._parameter_list
        self._m1_arena, m1 = self._arena(ps, FLOAT32)
        self._m2_arena, m2 = self._arena(ps, FLOAT32)
        for p, a, b in zip(ps, m1, m2):
            self._accumulators[p.name] = {self._moment1_acc_str: a, self._moment2_acc_str: b}
        masters = [p for p in ps if self._multi_precision and self._is_dtype_fp16_or_bf16(p.dtype)]
        if masters:
            self._master_arena, mw = self._arena(masters, FLOAT32)
            for p, w in zip(masters, mw):  # master = fp32 cast of the parameter (ref optimizer.py:640-666)
                a, b = p.prb, w.prb
                call("prb_cast", C.byref(a), C.byref(b), core._stream_handle(None))
                self._master_weights[p.name] = w
        # fill_value = beta (ref adamw.py:343-362): the first step divides by 1 - beta^1
        self._beta1_pow = Tensor.from_numpy(np.array([self._beta1], np.float32))
        self._beta2_pow = Tensor.from_numpy(np.array([self._beta2], np.float32))
        self._created = True

    def _find_master(self, p):
        return self._multi_precision and self._is_dtype_fp16_or_bf16(p.dtype)

    # ---- step
    def _params_grads(self):
        return [(p, p._reduced_grad or p.grad) for p in self._parameter_list
                if not p.stop_gradient and (p._reduced_grad is not None or p.grad is not None)]

    def step(self):
        """ref optimizer.py:1447-1504 -> _apply_optimize :1178-1221.

        The per-step host work is a cache lookup plus one C call per launch: the argument blocks of the
        fused tail are built once per distinct set of buffers and options (``_step_plans``)."""
        self._create_accumulators()
        pg = self._params_grads()
        if not pg:
            return
        found_inf = self._get_auxiliary_var("found_inf")
        loss_scale = self._get_auxiliary_var("loss_scale")  # set by GradScaler when the unscale is deferred to us
        clip = self._grad_clip
        dp0 = pg[0][0]._bucket[0]
        if dp0 is not None and dp0._sharded_armed:
            raise _lib.PRBError(3, "this backward was reduce-scattered (DataParallel.set_sharded): every rank holds the "
                                   "reduced gradient of its own arena ranges only; step through "
                                   "GroupShardedOptimizerStage2")
        if any(isinstance(g, SelectedRows) for _, g in pg):
            if loss_scale is not None:
                # the reference cannot either: check_finite_and_unscale_ has no SelectedRows kernel (legacy_ops.yaml:292-302)
                raise _lib.PRBError(4, "AMP unscaling of SelectedRows gradients is not supported")
            if clip is not None:
                # the global norm spans dense and sparse gradients: the reference-shaped clip (new gradient objects,
                # sparse norms from merged rows, clip.py:643-736), then one operator call per parameter
                pg = clip(pg)
            dense = [(p, g) for p, g in pg if not isinstance(g, SelectedRows)]
            self._sparse_step([(p, g) for p, g in pg if isinstance(g, SelectedRows)], found_inf,
                              advance_beta_pow=not dense)
            if clip is not None:
                if dense:
                    self._per_tensor_step(dense, found_inf, None, clipped=True)
                return
            pg = dense
            if not pg:
                return
        # everything a cached plan bakes in: buffers, flags and hyper-parameters (changing weight decay or the clip
        # norm in the middle of training builds a new plan instead of being ignored)
        key = (tuple(g.data_ptr for _, g in pg), id(clip), loss_scale.data_ptr if loss_scale is not None else 0,
               found_inf.data_ptr if found_inf is not None else 0, self._zero_grads_in_step,
               self._write_back_unscaled_grads, _clip_by_global_norm_using_mp_type(), self._hyper_key(),
               (clip.clip_norm, clip.auto_skip_clip) if clip is not None else None,
               pg[0][0]._bucket[0].pipelined if pg[0][0]._bucket[0] is not None else None, self._use_producer_pass1)
        plan = self._step_plans.get(key, _MISSING)
        if plan is _MISSING:
            if len(self._step_plans) > 16:
                self._step_plans.clear()
            plan = self._step_plans[key] = self._build_plan(pg, found_inf, loss_scale)
        if plan is None:  # mixed dtypes or per-parameter clip exclusions: the reference-shaped loop
            return self._per_tensor_step(pg, found_inf, loss_scale)
        lr = self._global_lr().prb
        lr_ref = C.byref(lr)
        stream = core._stream_handle(None)
        for entry in plan:
            dp, gi, args, ep = entry["dp"], entry["group"], entry["args"], entry["epilogue"]
            if dp is not None:
                if gi is None:
                    dp.wait_all_groups()
                else:
                    dp.wait_group(gi)
                if ep is not None and dp._epilogue_key == ep["key"] and dp._epilogue_armed:
                    # the backward that just ended ran with the epilogue: found_inf and the norm partials are there
                    args = entry["args_fused"]
                elif ep is not None or dp._epilogue_key is not None:
                    # (re)configure for the NEXT backward; this step still runs its own pass 1
                    if ep is not None and dp._configure_epilogue(ep["key"], loss_scale, found_inf, ep["partial_base"],
                                                                 _clip_by_global_norm_using_mp_type()):
                        if entry["args_fused"] is None:
                            entry["args_fused"] = self._tail_args(entry["pg"], found_inf, loss_scale,
                                                                  keep_beta_pow=False,
                                                                  pass1_partials=dp._epilogue_partials)
                    elif ep is None:
                        dp._configure_epilogue(None)
                    else:
                        entry["epilogue"] = None  # this reducer cannot carry it (NCCL / NVLS backend, arena mode)
            ev = self._launch_events
            if ev is not None:  # measurement hook: CUDA events right around each launch (after its stream wait)
                a, b = ev.pop()
                a.record()
                call("prb_dp_step_tail", args[0], args[1], args[2], lr_ref, *args[4:16], stream)
                b.record()
                self._launch_events_used.append((a, b))
            else:
                call("prb_dp_step_tail", args[0], args[1], args[2], lr_ref, *args[4:16], stream)

    def _hyper_key(self):
        return (float(self._beta1), float(self._beta2), float(self._epsilon), float(self._coeff()),
                id(getattr(self, "_lr_ratio", None)), id(getattr(self, "_apply_decay_param_fun", None)))

    def _tensor_attrs(self, pg):
        lr_ratio = np.ones(len(pg), np.float32)
        with_decay = np.ones(len(pg), np.uint8)
        return lr_ratio, with_decay

    def _build_plan(self, pg, found_inf, loss_scale):
        """-> list of plan entries {dp, group, args, pg, epilogue, args_fused}, or None for the slow path."""
        uniform = len({p.dtype for p, _ in pg}) == 1 and len({g.dtype for _, g in pg}) == 1 and \
            len({self._find_master(p) for p, _ in pg}) == 1
        clip = self._grad_clip
        if not uniform or (clip is not None and not all(p.need_clip is not False for p, _ in pg)):
            return None

        def entry(dp, gi, sub, fi, ls, keep, ep=None):
            return {"dp": dp, "group": gi, "pg": sub, "epilogue": ep, "args_fused": None,
                    "args": self._tail_args(sub, fi, ls, keep_beta_pow=keep)}

        # Gradients that sit in a DataParallel bucket arena: without a global dependency (no clip, no AMP) each
        # bucket's parameters are updated as soon as THAT bucket's allreduce is done, overlapping the allreduce of
        # the buckets behind it.  With a global dependency the producer of the buckets carries pass 1 (epilogue).
        dp = pg[0][0]._bucket[0]
        if dp is None:
            return [entry(None, None, pg, found_inf, loss_scale, False)]
        needs_pass1 = clip is not None or loss_scale is not None
        if dp.pipelined and not needs_pass1 and found_inf is None:
            by_group = {}
            for p, g in pg:
                by_group.setdefault(p._bucket[1], []).append((p, g))
            order = sorted(by_group)
            return [entry(dp, gi, by_group[gi], None, None, k != len(order) - 1) for k, gi in enumerate(order)]
        ep = None
        # AMP already unscaled by an explicit scaler.unscale_() (found_inf without loss_scale): the check and the norm
        # must see what that call left behind, so the producer cannot have done them
        if self._use_producer_pass1 and needs_pass1 and not (found_inf is not None and loss_scale is None):
            index = {id(p): i for i, p in enumerate(dp._parameters)}
            if len(pg) == len(index) and all(id(p) in index for p, _ in pg):
                chunk = lib.prb_grad_chunk_elems()
                base = np.zeros(len(pg), np.int32)
                at = 0
                for p, _ in pg:
                    base[index[id(p)]] = at
                    at += (p.numel + chunk - 1) // chunk
                ep = {"key": (id(self), loss_scale.data_ptr if loss_scale is not None else 0, clip is not None,
                              _clip_by_global_norm_using_mp_type()),
                      "partial_base": base if clip is not None else None}
        return [entry(dp, None, pg, found_inf, loss_scale, False, ep)]

    def _tail_args(self, pg, found_inf, loss_scale, keep_beta_pow, pass1_partials=None):
        """ctypes argument block of one prb_dp_step_tail call (kept alive by the plan that holds it)."""
        clip = self._grad_clip
        params = [p for p, _ in pg]
        grads = [g for _, g in pg]
        mp = self._find_master(params[0])
        cfg = TailConfig()
        cfg.beta1, cfg.beta2, cfg.epsilon = float(self._beta1), float(self._beta2), float(self._epsilon)
        cfg.coeff = float(self._coeff())
        cfg.use_adamw = int(self._use_adamw)
        cfg.multi_precision = int(mp)
        cfg.use_amp = int(loss_scale is not None)
        cfg.use_clip = int(clip is not None)
        cfg.clip_norm = float(clip.clip_norm) if clip else 0.0
        cfg.auto_skip_clip = int(clip.auto_skip_clip) if clip else 0
        cfg.use_mp_type = int(_clip_by_global_norm_using_mp_type())
        # p.grad holds the unscaled (never the clipped) gradient after a reference step
        cfg.write_back_grads = int(loss_scale is not None and self._write_back_unscaled_grads
                                   and not self._zero_grads_in_step)
        cfg.zero_grads = int(self._zero_grads_in_step)
        cfg.keep_beta_pow = int(keep_beta_pow)
        if pass1_partials is not None:
            cfg.pass1_done, cfg.pass1_partials = 1, pass1_partials.data_ptr
        lr_ratio, with_decay = self._tensor_attrs(pg)
        acc = [self._accumulators[p.name] for p in params]
        keep = {
            "cfg": cfg, "params": tensor_array(params), "grads": tensor_array(grads),
            "m1": tensor_array([a[self._moment1_acc_str] for a in acc]),
            "m2": tensor_array([a[self._moment2_acc_str] for a in acc]),
            "masters": tensor_array([self._master_weights[p.name] for p in params]) if mp else None,
            "lr_ratio": lr_ratio, "with_decay": with_decay, "b1": self._beta1_pow.prb, "b2": self._beta2_pow.prb,
            "ls": None, "fi": None, "gn": None, "cl": None,
        }
        if clip is not None:
            clip._ensure_outputs()
            keep["gn"], keep["cl"] = clip.last_global_norm.prb, clip.last_clip.prb
        if loss_scale is not None:
            keep["ls"] = loss_scale.prb
        if found_inf is not None:  # with use_amp=0 it is an input-only skip flag (unscale_ ran earlier)
            keep["fi"] = found_inf.prb
        ref = lambda k: C.byref(keep[k]) if keep[k] is not None else None  # noqa: E731
        args = (C.byref(cfg), keep["params"], keep["grads"], None, keep["m1"], keep["m2"], C.byref(keep["b1"]),
                C.byref(keep["b2"]), keep["masters"], len(params), lr_ratio.ctypes.data, with_decay.ctypes.data,
                ref("ls"), ref("fi"), ref("gn"), ref("cl"))
        return args + (keep,)  # the trailing dict only keeps the ctypes objects alive

    def _per_tensor_step(self, pg, found_inf, loss_scale, clipped=False):
        """Reference-shaped loop: one adam(w)_ call per parameter (ref adamw.py:445-465)."""
        if loss_scale is not None:
            by_dtype = {}
            for _, g in pg:
                by_dtype.setdefault(g.dtype, []).append(g)
            tmp = Tensor.zeros(1, BOOL)
            found_inf.zero_()
            for grads in by_dtype.values():  # one call per dtype list, flags OR-ed (ref grad_scaler.py:367-396)
                _C_ops.check_finite_and_unscale_(grads, loss_scale, tmp)
                _or_into(found_inf, tmp)
        if self._grad_clip is not None and not clipped:
            pg = self._grad_clip(pg)
        lr = self._global_lr()
        lr_ratio, with_decay = self._tensor_attrs(pg)
        for i, (p, g) in enumerate(pg):
            acc = self._accumulators[p.name]
            master = self._master_weights.get(p.name) if self._find_master(p) else None
            # the shared device accumulators advance once per step: with the last parameter
            self._append_optimize_op(p, g, lr, acc[self._moment1_acc_str], acc[self._moment2_acc_str], master, found_inf,
                                     float(lr_ratio[i]), bool(with_decay[i]), use_global_beta_pow=i != len(pg) - 1)

    def _sparse_step(self, pg, found_inf, advance_beta_pow):
        """Parameters whose gradient is a SelectedRows (embedding tables): one adam_ call each, which the YAML
        dispatches to adam_dense_param_sparse_grad (legacy_ops.yaml:52; ref adam.py:390-470).  They run BEFORE the
        dense parameters' fused launch, which is the one that advances the shared beta-pow accumulators."""
        if self._use_adamw:
            raise _lib.PRBError(4, "adamw_ has no SelectedRows kernel (legacy_ops.yaml:68-77); use Adam for sparse gradients")
        lr = self._global_lr()
        for i, (p, g) in enumerate(pg):
            acc = self._accumulators[p.name]
            master = self._master_weights.get(p.name) if self._find_master(p) else None
            last = advance_beta_pow and i == len(pg) - 1
            self._append_optimize_op(p, g, lr, acc[self._moment1_acc_str], acc[self._moment2_acc_str], master, found_inf,
                                     1.0, False, use_global_beta_pow=not last)

    # ---- housekeeping
    def clear_grad(self, set_to_zero=True):
        """ref optimizer.py:1333-1381 (one set_constant launch per gradient in the reference)."""
        for p in self._parameter_list:
            p.clear_gradient(set_to_zero)

    def state_dict(self):
        """Reference key layout (ref optimizer.py:294-327): accumulators by variable name, plus
        ``master_weights`` and ``LR_Scheduler``."""
        self._create_accumulators()
        sd = OrderedDict()
        b1 = self._beta1_pow.numpy()
        b2 = self._beta2_pow.numpy()
        for p in self._parameter_list:
            base = p.name + "_fp32_master_0" if self._find_master(p) else p.name
            acc = self._accumulators[p.name]
            sd[base + "_moment1_0"] = acc[self._moment1_acc_str].numpy()
            sd[base + "_moment2_0"] = acc[self._moment2_acc_str].numpy()
            sd[base + "_beta1_pow_acc_0"] = b1.copy()
            sd[base + "_beta2_pow_acc_0"] = b2.copy()
        if self._master_weights:
            sd["master_weights"] = {k: v.numpy() for k, v in self._master_weights.items()}
        if isinstance(self._learning_rate, LRScheduler):
            sd["LR_Scheduler"] = self._learning_rate.state_dict()
        return sd

    def set_state_dict(self, state_dict):
        """ref optimizer.py:329-409."""
        self._create_accumulators()
        if isinstance(self._learning_rate, LRScheduler) and "LR_Scheduler" in state_dict:
            self._learning_rate.set_state_dict(state_dict["LR_Scheduler"])
        for name, w in state_dict.get("master_weights", {}).items():
            if name in self._master_weights:
                self._master_weights[name].copy_from_numpy(np.asarray(w, np.float32))
        first = True
        for p in self._parameter_list:
            base = p.name + "_fp32_master_0" if self._find_master(p) else p.name
            acc = self._accumulators[p.name]
            for key, t in ((base + "_moment1_0", acc[self._moment1_acc_str]), (base + "_moment2_0", acc[self._moment2_acc_str])):
                if key not in state_dict:
                    raise KeyError("optimizer variable %s not found" % key)
                t.copy_from_numpy(np.asarray(state_dict[key], np.float32).reshape(-1))
            if first:
                self._beta1_pow.copy_from_numpy(np.asarray(state_dict[base + "_beta1_pow_acc_0"], np.float32).reshape(-1))
                self._beta2_pow.copy_from_numpy(np.asarray(state_dict[base + "_beta2_pow_acc_0"], np.float32).reshape(-1))
                first = False

    set_dict = set_state_dict


def _or_into(dst: Tensor, src: Tensor):
    """found_inf = bitwise_or(found_inf, tmp) on the device (ref grad_scaler.py:373-375)."""
    call("prb_flag_or", dst.data_ptr, src.data_ptr, core._stream_handle(None))
