"""CPU oracle for the data-parallel step tail -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module; the product package
``paddle_ray_b200`` never does.  Arithmetic lives in ``dp_tail_oracle.c`` (built to
``liboracle.so`` by ``oracle/Makefile``); this file adds the integer/host logic of
the path, restated from the reference (paths relative to /root/reference):

* ``assign_group_by_size``  <- paddle/fluid/distributed/collective/reducer.cc:45-156
* ``ReducerModel``          <- reducer.cc:669-728 (PrepareForBackward), :774-901
                               (MarkVarReady), :903-930 (MarkGroupReady), :1013-1035
* ``global_norm_clip``      <- python/paddle/nn/clip.py:643-736 (_dygraph_clip)
* ``dp_step_tail``          <- composition of the above in the order of SURVEY.md 3.2/3.3

Parity status: pinned against the reference's own test oracles and literal vectors
(see tests/golden/make_golden.py); summation order of the norm and of the
allreduce is unpinned by the reference and compared with tolerance.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

BOOL, INT32, INT64, F32, F64, F16, BF16 = 1, 7, 9, 10, 11, 15, 16
SIZEOF = {BOOL: 1, INT32: 4, INT64: 8, F32: 4, F64: 8, F16: 2, BF16: 2}
NP_STORAGE = {F32: np.float32, F64: np.float64, F16: np.uint16, BF16: np.uint16}


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "dp_tail_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"], env={**os.environ, "CC": "gcc"})
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.orc_f32_to_bf16.restype = C.c_uint16
        L.orc_f32_to_bf16.argtypes = [C.c_float]
        L.orc_bf16_to_f32.restype = C.c_float
        L.orc_bf16_to_f32.argtypes = [C.c_uint16]
        L.orc_f32_to_f16.restype = C.c_uint16
        L.orc_f32_to_f16.argtypes = [C.c_float]
        L.orc_f16_to_f32.restype = C.c_float
        L.orc_f16_to_f32.argtypes = [C.c_uint16]
        L.orc_convert_from_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        L.orc_convert_to_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        L.orc_adam_f32.argtypes = (
            [C.c_int, C.c_int, C.c_int64] + [C.c_void_p] * 5 + [C.c_float] * 6 + [C.c_int] + [C.c_float] * 2 + [C.c_int]
        )
        L.orc_adam_f64.argtypes = [C.c_int64] + [C.c_void_p] * 4 + [C.c_double] * 6 + [C.c_int] + [C.c_double] * 2
        L.orc_update_beta_pow_f32.argtypes = [C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_scale.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_double, C.c_float, C.c_int]
        L.orc_concat.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_split.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_check_finite_and_unscale.restype = C.c_int
        L.orc_check_finite_and_unscale.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_float]
        L.orc_squared_l2_norm.restype = C.c_float
        L.orc_squared_l2_norm.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_squared_l2_norm_mp.restype = C.c_float
        L.orc_squared_l2_norm_mp.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_clip_mul.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_float]
        L.orc_update_loss_scaling.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float]
        L.orc_cpu_adamw.argtypes = [C.c_int64] + [C.c_void_p] * 4 + [C.c_float] * 8 + [C.c_int]
        L.orc_cpu_scale_f32.argtypes = [C.c_int64, C.c_void_p, C.c_float, C.c_int]
        L.orc_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


# ---------------------------------------------------------------- conversions
def from_f32(x: np.ndarray, dtype: int) -> np.ndarray:
    """float32 array -> storage array of ``dtype`` (uint16 bit patterns for bf16/f16)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    if dtype == F32:
        return x.copy()
    if dtype == F64:
        return x.astype(np.float64)
    out = np.empty(x.shape, dtype=np.uint16)
    lib().orc_convert_from_f32(_p(x), _p(out), x.size, dtype)
    return out


def to_f32(x: np.ndarray, dtype: int) -> np.ndarray:
    if dtype == F32:
        return np.ascontiguousarray(x, dtype=np.float32).copy()
    if dtype == F64:
        return x.astype(np.float32)
    x = np.ascontiguousarray(x, dtype=np.uint16)
    out = np.empty(x.shape, dtype=np.float32)
    lib().orc_convert_to_f32(_p(x), _p(out), x.size, dtype)
    return out


# ---------------------------------------------------------------- adam family
def adam(param, grad, mom1, mom2, master, *, dtype_param, dtype_grad, lr, lr_ratio=1.0, beta1, beta2, epsilon,
         coeff=0.0, use_decay=False, beta1_pow, beta2_pow, fma=False):
    """In-place Adam/AdamW update (GPU semantics; adamw_kernel.cu:32-79, adam_kernel.cu:31-72)."""
    n = param.size
    if dtype_param == F64:
        lib().orc_adam_f64(n, _p(param), _p(grad), _p(mom1), _p(mom2), lr, lr_ratio, beta1, beta2, epsilon, coeff,
                           int(use_decay), beta1_pow, beta2_pow)
        return
    lib().orc_adam_f32(dtype_param, dtype_grad, n, _p(param), _p(grad), _p(mom1), _p(mom2), _p(master),
                       lr, lr_ratio, beta1, beta2, epsilon, coeff, int(use_decay), beta1_pow, beta2_pow, int(fma))


def update_beta_pow(beta1, beta2, b1p: np.ndarray, b2p: np.ndarray):
    lib().orc_update_beta_pow_f32(beta1, beta2, _p(b1p), _p(b2p))


# ---------------------------------------------------------------- bucket ops
def scale(x: np.ndarray, dtype: int, scale_d: float, bias: float = 0.0, bias_after_scale: bool = False) -> np.ndarray:
    out = np.empty_like(x)
    lib().orc_scale(dtype, x.size, _p(x), _p(out), float(scale_d), float(bias), int(bias_after_scale))
    return out


def concat(arrays, out=None) -> np.ndarray:
    """Dense byte concat of same-dtype 1-D arrays (concat_and_split_functor.cc:52-66)."""
    arrays = [np.ascontiguousarray(a) for a in arrays]
    total = sum(a.size for a in arrays)
    if out is None:
        out = np.empty(total, dtype=arrays[0].dtype)
    ptrs = (C.c_void_p * len(arrays))(*[a.ctypes.data for a in arrays])
    nb = np.array([a.nbytes for a in arrays], dtype=np.int64)
    lib().orc_concat(len(arrays), ptrs, _p(nb), _p(out))
    return out


def split(bucket: np.ndarray, numels, outs=None) -> list:
    if outs is None:
        outs = [np.empty(int(n), dtype=bucket.dtype) for n in numels]
    ptrs = (C.c_void_p * len(outs))(*[o.ctypes.data for o in outs])
    nb = np.array([o.nbytes for o in outs], dtype=np.int64)
    lib().orc_split(len(outs), _p(bucket), _p(nb), ptrs)
    return outs


# ---------------------------------------------------------------- amp / clip
def check_finite_and_unscale(xs, dtype: int, scale_v: float) -> bool:
    """In place on each array of ``xs``; returns found_inf (amp_kernel.cu:27-82)."""
    ptrs = (C.c_void_p * len(xs))(*[x.ctypes.data for x in xs])
    ns = np.array([x.size for x in xs], dtype=np.int64)
    return bool(lib().orc_check_finite_and_unscale(dtype, len(xs), ptrs, _p(ns), float(scale_v)))


def squared_l2_norm(x: np.ndarray, dtype: int, mp: bool = False):
    """Returns (value as the op would type it, as float; fp64 reference sum)."""
    f64 = C.c_double(0.0)
    fn = lib().orc_squared_l2_norm_mp if mp else lib().orc_squared_l2_norm
    v = fn(dtype, x.size, _p(x), C.byref(f64))
    return float(v), float(f64.value)


def clip_mul(g: np.ndarray, dtype: int, clip: float):
    lib().orc_clip_mul(dtype, g.size, _p(g), float(clip))


def global_norm_clip(grads, dtypes, clip_norm: float, auto_skip_clip: bool = False, mp: bool = False):
    """ClipGradByGlobalNorm._dygraph_clip (python/paddle/nn/clip.py:643-736), in place.

    Per-tensor squared norms are grouped into a 16-bit list and an fp32 list, each
    summed (add_n, float accumulation; the 16-bit partial sum is itself a T-typed
    tensor), cast to fp32, added, sqrt'ed; ``clip = max_norm / max(norm, max_norm)``.
    Returns (global_norm, clip_coeff or None when skipped).  The sums are
    accumulated in float64 here and narrowed once: the reference leaves the
    order unspecified, so callers compare with tolerance.
    """
    s16, s32 = [], []
    dt16 = None
    for g, dt in zip(grads, dtypes):
        v, v64 = squared_l2_norm(g, dt, mp=mp)
        if dt in (F16, BF16) and not mp:
            # the op's T-typed output; the order-free fp64 sum is narrowed once (a sequential float sum of
            # millions of T-rounded squares stagnates and is NOT what a tree reduction on the GPU produces)
            s16.append(float(to_f32(from_f32(np.array([v64], np.float32), dt), dt)[0]))
            dt16 = dt
        else:
            s32.append(v64 if dt == F32 or mp else v)
    parts = []
    if s16:
        acc = np.float32(0)
        for v in s16:
            acc = np.float32(acc + np.float32(v))
        # add_n of T-typed scalars yields a T-typed scalar (accumulated in float), then astype(fp32)
        acc = to_f32(from_f32(np.array([acc], np.float32), dt16), dt16)[0]
        parts.append(np.float32(acc))
    if s32:
        parts.append(np.float32(np.sum(np.array(s32, dtype=np.float64))))
    if not parts:
        return None, None
    gsum = np.float32(0)
    for p_ in parts:
        gsum = np.float32(gsum + p_)
    norm = np.float32(np.sqrt(gsum))
    mx = np.float32(clip_norm)
    if not auto_skip_clip:
        clip = np.float32(mx / np.maximum(norm, mx))
    elif norm > mx:
        clip = np.float32(mx / norm)
    else:
        return float(norm), None
    for g, dt in zip(grads, dtypes):
        clip_mul(g, dt, float(clip))
    return float(norm), float(clip)


def update_loss_scaling(found_inf, loss_scaling, good, bad, incr_every_n_steps, decr_every_n_nan_or_inf, incr_ratio,
                        decr_ratio):
    ls = np.array([loss_scaling], dtype=np.float32)
    g = np.array([good], dtype=np.int32)
    b = np.array([bad], dtype=np.int32)
    lib().orc_update_loss_scaling(int(found_inf), _p(ls), _p(g), _p(b), incr_every_n_steps, decr_every_n_nan_or_inf,
                                  incr_ratio, decr_ratio)
    return float(ls[0]), int(g[0]), int(b[0])


# ---------------------------------------------------------------- CPU baseline (timed, not the parity target)
def cpu_adamw(param, grad, mom1, mom2, *, lr, lr_ratio, beta1, beta2, epsilon, coeff, beta1_pow, beta2_pow, nthreads):
    lib().orc_cpu_adamw(param.size, _p(param), _p(grad), _p(mom1), _p(mom2), lr, lr_ratio, beta1, beta2, epsilon,
                        coeff, beta1_pow, beta2_pow, nthreads)


def cpu_scale_f32(x, scale_v, nthreads):
    lib().orc_cpu_scale_f32(x.size, _p(x), scale_v, nthreads)


def max_threads() -> int:
    return int(lib().orc_max_threads())


# ---------------------------------------------------------------- bucket assignment
def assign_group_by_size(numels, dtypes, is_sparse, group_size_limits, tensor_indices=()):
    """Eager_AssignGroupBySize (reducer.cc:45-156).

    ``numels[i]``/``dtypes[i]`` describe ``tensors[i]`` (the i-th tensor *as passed*);
    with ``tensor_indices`` the i-th tensor is recorded under index
    ``tensor_indices[i]`` (and its sparsity looked up there), exactly as the
    reference does (:89-96).  Leftover open groups are appended in dtype-enum
    order (std::map), and groups are sorted by first index only when no
    ``tensor_indices`` were given (:147-154).
    """
    n = len(numels)
    if len(is_sparse) != n:
        raise ValueError("tensors len must be equal to is_sparse_gradient len")
    ti = list(tensor_indices)
    if ti and sorted(ti) != list(range(len(ti))):
        raise ValueError("tensor_indices must be a permutation")
    res = []
    limit_index = {}
    next_group = {}
    for i in range(n):
        real = ti[i] if ti else i
        if is_sparse[real]:
            res.append([real])
            continue
        dt = dtypes[i]
        grp = next_group.setdefault(dt, [[], 0])
        grp[0].append(real)
        grp[1] += SIZEOF[dt] * numels[i]
        limit_index.setdefault(dt, 0)
        if grp[1] >= group_size_limits[limit_index[dt]]:
            res.append(grp[0])
            next_group[dt] = [[], 0]
            limit_index[dt] = min(limit_index[dt] + 1, len(group_size_limits) - 1)
    for dt in sorted(next_group):
        if next_group[dt][0]:
            res.append(next_group[dt][0])
    if not ti:
        res.sort(key=lambda g: g[0])
    return res


# ---------------------------------------------------------------- reducer state machine model
class ReducerModel:
    """Host state machine of EagerReducer, dense groups only (reducer.cc:669-930,1013-1035).

    ``launched`` records the order in which groups are handed to
    FusedAllReduceSchedule; ``finalized`` flips when next_group_ reaches the end.
    """

    def __init__(self, group_indices, ntensors):
        self.groups = [list(g) for g in group_indices]
        self.locator = {}
        for gi, g in enumerate(self.groups):
            for si, v in enumerate(g):
                self.locator[v] = (gi, si)
        self.ntensors = ntensors
        self.need_finalize = False
        self.grad_need_hooks = False
        self.prepare_for_backward()

    def prepare_for_backward(self):
        if self.need_finalize:
            raise RuntimeError("previous backward was not finalized")  # reducer.cc:683-698
        self.grad_need_hooks = True
        self.next_group = 0
        self.pending = [len(g) for g in self.groups]
        self.ready = [False] * self.ntensors
        self.launched = []
        self.finalized = False

    def mark_var_ready(self, v):
        if not self.grad_need_hooks:  # no_sync / after finalize: only re-alias (reducer.cc:740-758)
            return
        if self.ready[v]:
            raise RuntimeError("parameter %d has been ready before" % v)  # :779-811
        self.ready[v] = True
        self.need_finalize = True
        gi, _ = self.locator[v]
        self.pending[gi] -= 1
        if self.pending[gi] == 0:
            self._mark_group_ready(gi)
        if self.next_group == len(self.groups):
            self.need_finalize = False
            self.grad_need_hooks = False
            self.finalized = True

    def _mark_group_ready(self, gi):
        if gi < self.next_group:
            raise RuntimeError("group order violated")  # :906-914
        if gi > self.next_group:
            return  # deferred (:916-919)
        while self.next_group < len(self.groups) and self.pending[self.next_group] == 0:
            self.launched.append(self.next_group)
            self.next_group += 1


# ---------------------------------------------------------------- sparse (SelectedRows) gradients
def merge_rows(rows, values, dtype, width=None):
    """scatter::MergeAdd (paddle/phi/kernels/funcs/selected_rows_functor.cu, used by
    selected_rows/gpu/adam_kernel.cu:180-206): unique rows in ascending order, the value rows of duplicates added.
    The reference adds with atomics (order unspecified); here: fp32 accumulation in input order, rounded once to T
    -- identical whenever a row occurs at most twice."""
    rows = np.asarray(rows, np.int64)
    values = np.ascontiguousarray(values)
    if width is None:
        width = values.size // max(rows.size, 1)
    v32 = to_f32(values.reshape(-1), dtype).reshape(rows.size, width)
    urows = np.unique(rows)
    acc = np.zeros((urows.size, width), np.float32)
    pos = np.searchsorted(urows, rows)
    for k in range(rows.size):            # input order
        acc[pos[k]] = acc[pos[k]] + v32[k]
    return urows, from_f32(acc.reshape(-1), dtype).reshape(urows.size, width)


def sparse_adam(param, rows, values, mom1, mom2, master, *, dtype_param, lr, beta1, beta2, epsilon, beta1_pow,
                beta2_pow, lazy_mode):
    """adam with a SelectedRows gradient (selected_rows/gpu/adam_kernel.cu:41-93): per element the dense Adam
    expression with g = the merged row's value, or 0 where the row is absent; lazy_mode leaves absent rows untouched.
    ``param`` / moments / master are 2-D [height, width] arrays updated in place."""
    width = param.shape[1]
    urows, merged = merge_rows(rows, values, dtype_param, width)
    kw = dict(dtype_param=dtype_param, dtype_grad=dtype_param, lr=lr, beta1=beta1, beta2=beta2, epsilon=epsilon,
              beta1_pow=beta1_pow, beta2_pow=beta2_pow)
    if lazy_mode:
        for k, r in enumerate(urows):
            adam(param[r], merged[k], mom1[r], mom2[r], None if master is None else master[r], **kw)
        return
    dense = from_f32(np.zeros(param.size, np.float32), dtype_param).reshape(param.shape[0], width)
    dense[urows] = merged
    adam(param.reshape(-1), dense.reshape(-1), mom1.reshape(-1), mom2.reshape(-1),
         None if master is None else master.reshape(-1), **kw)


def sparse_allreduce(per_rank, dtype):
    """EagerReducer::AllReduceSparse (reducer.cc:1072-1277): every rank's values are scaled by 1/nranks (:1075-1076),
    then rows and values of all ranks are concatenated in rank order (all-gather, or rank-by-rank broadcasts + concat
    when the ranks hold different numbers of rows: same result).  ``per_rank``: list of (rows, values[nrows, width])."""
    n = len(per_rank)
    rows = np.concatenate([np.asarray(r, np.int64) for r, _ in per_rank])
    width = max(np.asarray(v).shape[1] for _, v in per_rank)
    parts = [scale(np.ascontiguousarray(v).reshape(-1), dtype, 1.0 / n).reshape(-1, width)
             for _, v in per_rank if np.asarray(v).size]
    vals = np.concatenate(parts, axis=0) if parts else np.zeros((0, width), np.asarray(per_rank[0][1]).dtype)
    return rows, vals
