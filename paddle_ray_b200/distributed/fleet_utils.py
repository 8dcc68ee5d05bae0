"""Fleet-style gradient sync: coalesce -> scale -> allreduce -> split, NOT overlapped with backward.

ref: python/paddle/distributed/fleet/utils/hybrid_parallel_util.py:73-111 (_apply_collective_grads_eager),
     :199-216 (fused_allreduce_gradients[_with_group]); python/paddle/distributed/parallel.py:73-116
     (_coalesce_tensors / _split_tensors), :136-153 (build_groups, 128 MiB buckets).

Same kernels as the reducer path (multi-tensor concat / scale / split), different host schedule: every bucket is
synchronised on the calc stream after backward has finished.  ``param.main_grad`` (an fp32 gradient kept next to
a 16-bit parameter) takes precedence over ``param.grad`` like in the reference (:83-88).
"""
from __future__ import annotations

from collections import OrderedDict

from .. import _C_ops, _lib
from ..core import Tensor
from . import parallel as _parallel


def build_groups(vars, group_size):
    """ref parallel.py:136-153: consecutive same-dtype tensors; a new group starts once the running byte count has
    reached ``group_size`` (checked BEFORE adding, so a group may exceed it by one tensor) or the dtype changes."""
    group_idx = 0
    memory_counter = 0
    var_groups = OrderedDict()
    dtype = vars[0].dtype
    for var in vars:
        nbytes = var.numel * _lib.DTYPE_SIZE[var.dtype]
        if memory_counter < group_size and dtype == var.dtype:
            memory_counter += nbytes
        else:
            memory_counter = nbytes
            dtype = var.dtype
            group_idx += 1
        var_groups.setdefault(group_idx, []).append(var)
    return list(var_groups.values())


_coalesced_cache: dict = {}


def _coalesced_buffer(grad_vars) -> Tensor:
    """One persistent flat buffer per distinct group (the reference concatenates into a fresh tensor every step)."""
    key = tuple(g.data_ptr for g in grad_vars)
    buf = _coalesced_cache.get(key)
    if buf is None:
        if len(_coalesced_cache) > 256:
            _coalesced_cache.clear()
        buf = _coalesced_cache[key] = Tensor(sum(g.numel for g in grad_vars), grad_vars[0].dtype, grad_vars[0].device)
    return buf


def fused_allreduce_gradients_with_group(parameter_list, group=None, bucket_size=128 * 1024 * 1024, scale=None,
                                         stream=None):
    """ref hybrid_parallel_util.py:199-209 -> _apply_collective_grads_eager :73-111."""
    grad_vars, seen = [], set()
    for param in parameter_list:
        g_var = None
        if not param.stop_gradient and param.grad is not None:
            g_var = param.grad
        if not param.stop_gradient and getattr(param, "main_grad", None) is not None:
            g_var = param.main_grad
        if g_var is not None:
            assert id(g_var) not in seen
            seen.add(id(g_var))
            grad_vars.append(g_var)
    if not grad_vars:
        return
    group = group if group is not None else _parallel.get_group()
    nranks = group.nranks if group is not None else 1
    scale = 1.0 / nranks if scale is None else scale
    scale = None if scale == 1.0 else scale
    groups = build_groups(grad_vars, bucket_size)
    coalesced = []
    for gv in groups:
        buf = _coalesced_buffer(gv)
        o = buf.prb
        _lib.call("prb_concat", _lib.tensor_array(gv), len(gv), _lib.C.byref(o), _parallel.core._stream_handle(stream))
        if scale is not None:
            _C_ops.scale_(buf, scale, 0.0, True, stream=stream)   # Tensor.scale_(scale): scale * x + 0
        if group is not None and nranks > 1:
            group.all_reduce(buf, "sum", stream=stream)
        coalesced.append((buf, gv))
    for buf, gv in coalesced:
        _C_ops.split_(buf, gv, stream=stream)


def fused_allreduce_gradients(parameter_list, hcg=None):
    """ref hybrid_parallel_util.py:212-216; ``hcg`` may expose get_data_parallel_group()."""
    group = None if hcg is None else hcg.get_data_parallel_group()
    fused_allreduce_gradients_with_group(parameter_list, group)
