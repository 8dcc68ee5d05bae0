"""Operator entry points with the names and argument order of the reference's generated
``paddle._C_ops`` (from paddle/phi/api/yaml/legacy_ops.yaml), each a thin call into the C ABI.

ref: adamw_ :68-77, adam_ :45-55, merged_adam_ :1090-1099, fused_adam_ :718-727,
     check_finite_and_unscale_ :292-302, squared_l2_norm :1504-1511, update_loss_scaling_ :1708-1717,
     scale ops.yaml:1251-1262.

All ops are in place where the reference's YAML says ``inplace``, are enqueued on
``core.current_stream()`` and never synchronise.
"""
from __future__ import annotations

import ctypes as C

from . import core
from ._lib import BOOL, FLOAT32, call, tensor_array
from .core import Tensor


def _ref(t):
    if t is None:
        return None
    p = t.prb
    return C.byref(p), p


def _opt(t):
    """optional Tensor -> (ctypes arg, keepalive)"""
    if t is None:
        return None, None
    p = t.prb
    return C.byref(p), p


def _stream(stream):
    return core._stream_handle(stream)


def adamw_(param, grad, learning_rate, moment1, moment2, beta1_pow, beta2_pow, master_param, skip_update, beta1, beta2,
           epsilon, lr_ratio, coeff, with_decay, lazy_mode=False, min_row_size_to_use_multithread=1000,
           multi_precision=False, use_global_beta_pow=False, stream=None):
    ts = [param, grad, learning_rate, moment1, moment2, beta1_pow, beta2_pow]
    ps = [t.prb for t in ts]
    mp, mk = _opt(master_param)
    sp, sk = _opt(skip_update)
    call("prb_adamw", *[C.byref(p) for p in ps], mp, sp, float(beta1), float(beta2), float(epsilon), float(lr_ratio),
         float(coeff), int(bool(with_decay)), int(bool(lazy_mode)), int(min_row_size_to_use_multithread),
         int(bool(multi_precision)), int(bool(use_global_beta_pow)), C.byref(ps[0]), C.byref(ps[3]), C.byref(ps[4]),
         C.byref(ps[5]), C.byref(ps[6]), mp, _stream(stream))
    return param, moment1, moment2, beta1_pow, beta2_pow, master_param


def adam_(param, grad, learning_rate, moment1, moment2, beta1_pow, beta2_pow, master_param, skip_update, beta1, beta2,
          epsilon, lazy_mode=False, min_row_size_to_use_multithread=1000, multi_precision=False,
          use_global_beta_pow=False, stream=None):
    if isinstance(grad, core.SelectedRows):
        # the YAML's kernel dispatch on the gradient's type: adam_dense_param_sparse_grad (legacy_ops.yaml:52)
        ts = [param, learning_rate, moment1, moment2, beta1_pow, beta2_pow]
        ps = [t.prb for t in ts]
        sr, keep = grad.c_struct()
        mp, mk = _opt(master_param)
        sp, sk = _opt(skip_update)
        call("prb_adam_dense_param_sparse_grad", C.byref(ps[0]), C.byref(sr), *[C.byref(p) for p in ps[1:]], mp, sp,
             float(beta1), float(beta2), float(epsilon), int(bool(lazy_mode)), int(min_row_size_to_use_multithread),
             int(bool(multi_precision)), int(bool(use_global_beta_pow)), C.byref(ps[0]), C.byref(ps[2]), C.byref(ps[3]),
             C.byref(ps[4]), C.byref(ps[5]), mp, _stream(stream))
        return param, moment1, moment2, beta1_pow, beta2_pow, master_param
    ts = [param, grad, learning_rate, moment1, moment2, beta1_pow, beta2_pow]
    ps = [t.prb for t in ts]
    mp, mk = _opt(master_param)
    sp, sk = _opt(skip_update)
    call("prb_adam", *[C.byref(p) for p in ps], mp, sp, float(beta1), float(beta2), float(epsilon),
         int(bool(lazy_mode)), int(min_row_size_to_use_multithread), int(bool(multi_precision)),
         int(bool(use_global_beta_pow)), C.byref(ps[0]), C.byref(ps[3]), C.byref(ps[4]), C.byref(ps[5]), C.byref(ps[6]),
         mp, _stream(stream))
    return param, moment1, moment2, beta1_pow, beta2_pow, master_param


def merge_selected_rows(x, stream=None):
    """-> a new SelectedRows: unique rows ascending, duplicates' value rows added (legacy_ops.yaml:1082-1088)."""
    import numpy as np
    sr, keep = x.c_struct()
    rows = np.zeros(max(x.rows.size, 1), np.int64)
    n = C.c_int64()
    value = Tensor(max(x.value.numel, 1), x.value.dtype, x.value.device)
    v = value.prb
    call("prb_merge_selected_rows", C.byref(sr), rows.ctypes.data, C.byref(n), C.byref(v), _stream(stream))
    width = x.width
    return core.SelectedRows(rows[:n.value].copy(), x.height, value.view(0, n.value * width))


def multiply_by_clip_(xs, clip, stream=None):
    """In place on the list: x = T(MT(x) * MT(T(clip)))  (the multiply of ClipGradByGlobalNorm, clip.py:715-727)."""
    c = clip.prb
    call("prb_multiply_by_clip", tensor_array(xs), len(xs), C.byref(c), _stream(stream))
    return xs


def merged_adam_(param, grad, learning_rate, moment1, moment2, beta1_pow, beta2_pow, master_param, beta1, beta2,
                 epsilon, multi_precision=False, use_global_beta_pow=False, stream=None):
    n = len(param)
    arrs = [tensor_array(x) for x in (param, grad, learning_rate, moment1, moment2, beta1_pow, beta2_pow)]
    ma = tensor_array(master_param) if master_param else None
    call("prb_merged_adam", *arrs, ma, n, float(beta1), float(beta2), float(epsilon), int(bool(multi_precision)),
         int(bool(use_global_beta_pow)), arrs[0], arrs[3], arrs[4], arrs[5], arrs[6], ma, _stream(stream))
    return param, moment1, moment2, beta1_pow, beta2_pow, master_param


def fused_adam_(params, grads, learning_rate, moments1, moments2, beta1_pows, beta2_pows, master_params, skip_update,
                beta1, beta2, epsilon, chunk_size, weight_decay, use_adamw, multi_precision=False,
                use_global_beta_pow=False, stream=None):
    n = len(params)
    arrs = [tensor_array(x) for x in (params, grads)]
    lr = learning_rate.prb
    rest = [tensor_array(x) for x in (moments1, moments2, beta1_pows, beta2_pows)]
    ma = tensor_array(master_params) if master_params else None
    sp, sk = _opt(skip_update)
    call("prb_fused_adam", arrs[0], arrs[1], C.byref(lr), *rest, ma, n, sp, float(beta1), float(beta2), float(epsilon),
         int(chunk_size), float(weight_decay), int(bool(use_adamw)), int(bool(multi_precision)),
         int(bool(use_global_beta_pow)), arrs[0], rest[0], rest[1], rest[2], rest[3], ma, _stream(stream))
    return params, moments1, moments2, beta1_pows, beta2_pows, master_params


def check_finite_and_unscale_(x, scale, found_infinite=None, stream=None):
    """In place on the list ``x``; returns (x, found_infinite)."""
    if found_infinite is None:
        found_infinite = Tensor(1, BOOL)
    xa = tensor_array(x)
    s, f = scale.prb, found_infinite.prb
    call("prb_check_finite_and_unscale", xa, len(x), C.byref(s), xa, C.byref(f), _stream(stream))
    return x, found_infinite


def update_loss_scaling_(x, found_infinite, prev_loss_scaling, in_good_steps, in_bad_steps, incr_every_n_steps,
                         decr_every_n_nan_or_inf, incr_ratio, decr_ratio, stop_update=False, stream=None):
    xa = tensor_array(x) if x else None
    f, p, g, b = found_infinite.prb, prev_loss_scaling.prb, in_good_steps.prb, in_bad_steps.prb
    call("prb_update_loss_scaling", xa, len(x) if x else 0, C.byref(f), C.byref(p), C.byref(g), C.byref(b),
         int(incr_every_n_steps), int(decr_every_n_nan_or_inf), float(incr_ratio), float(decr_ratio),
         int(bool(stop_update)), xa, C.byref(p), C.byref(g), C.byref(b), _stream(stream))
    return x, prev_loss_scaling, in_good_steps, in_bad_steps


def squared_l2_norm(x, stream=None):
    out = Tensor(1, x.dtype, x.device)
    a, b = x.prb, out.prb
    call("prb_squared_l2_norm", C.byref(a), C.byref(b), _stream(stream))
    return out


def scale(x, scale, bias=0.0, bias_after_scale=True, stream=None):
    out = Tensor(x.numel, x.dtype, x.device)
    a, b = x.prb, out.prb
    call("prb_scale", C.byref(a), float(scale), float(bias), int(bool(bias_after_scale)), C.byref(b), _stream(stream))
    return out


def scale_(x, scale, bias=0.0, bias_after_scale=True, stream=None):
    a = x.prb
    call("prb_scale", C.byref(a), float(scale), float(bias), int(bool(bias_after_scale)), C.byref(a), _stream(stream))
    return x


def concat(xs, stream=None):
    """Dense 1-D concat of same-dtype tensors (EagerGroup::ConcatTensors)."""
    total = sum(t.numel for t in xs)
    out = Tensor(total, xs[0].dtype, xs[0].device)
    o = out.prb
    call("prb_concat", tensor_array(xs), len(xs), C.byref(o), _stream(stream))
    return out


def split_(x, outs, stream=None):
    """Scatter ``x`` back into the tensors of ``outs`` (EagerGroup::SplitTensors)."""
    a = x.prb
    call("prb_split", C.byref(a), tensor_array(outs), len(outs), _stream(stream))
    return outs


def clip_grad_by_global_norm_(grads, clip_norm, auto_skip_clip=False, use_mp_type=False, stream=None, norm=None,
                              clip=None):
    """Multi-tensor body of ClipGradByGlobalNorm._dygraph_clip; returns (global_norm, clip) device scalars."""
    norm = Tensor(1, FLOAT32, grads[0].device) if norm is None else norm
    clip = Tensor(1, FLOAT32, grads[0].device) if clip is None else clip
    n, c = norm.prb, clip.prb
    call("prb_clip_grad_by_global_norm", tensor_array(grads), len(grads), float(clip_norm), int(bool(auto_skip_clip)),
         int(bool(use_mp_type)), C.byref(n), C.byref(c), _stream(stream))
    return norm, clip
