"""``init_parallel_env`` / ``DataParallel`` -- the host side of the gradient-bucket path.

ref: python/paddle/distributed/parallel.py:202-561 (DataParallel), :921-1211 (init_parallel_env),
     paddle/fluid/pybind/distributed_py.cc:64-77,1285-1315 (eager_assign_group_by_size, EagerReducer).

The reference wraps a ``Layer`` and hangs reduce hooks on the autograd engine.  Model code and
autograd are out of scope here, so ``DataParallel`` wraps the *parameter list*; the producer of
gradients calls ``mark_var_ready(i)`` (the hook equivalent) as each gradient lands.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile
import warnings

import numpy as np

from .. import _lib, core
from .._lib import call, lib, tensor_array
from ..core import Tensor
from .store import FileStore, TCPStore


class ParallelEnv:
    """Rank / world size / endpoints from the environment.

    Honours the reference's names (PADDLE_TRAINER_ID, PADDLE_TRAINERS_NUM, PADDLE_MASTER,
    FLAGS_selected_gpus; parallel.py:1030-1083) and, when those are absent, the torchrun names
    (RANK, WORLD_SIZE, LOCAL_RANK, MASTER_ADDR, MASTER_PORT).
    """

    def __init__(self):
        env = os.environ
        self.rank = int(env.get("PADDLE_TRAINER_ID", env.get("RANK", "0")))
        self.world_size = int(env.get("PADDLE_TRAINERS_NUM", env.get("WORLD_SIZE", "1")))
        sel = env.get("FLAGS_selected_gpus")
        self.device_id = int(sel.split(",")[0]) if sel else int(env.get("LOCAL_RANK", str(self.rank)))
        master = env.get("PADDLE_MASTER")
        if master:
            host, port = master.rsplit(":", 1)
        else:
            host, port = env.get("MASTER_ADDR", "127.0.0.1"), env.get("MASTER_PORT", "6170")
        self.master_addr, self.master_port = host, int(port)
        self.under_torchrun = "TORCHELASTIC_RUN_ID" in env or ("LOCAL_RANK" in env and "PADDLE_TRAINER_ID" not in env)

    @property
    def nranks(self):
        return self.world_size

    @property
    def local_rank(self):
        return self.rank


class Group:
    """A communicator handle (``paddle.distributed.collective.Group`` + its ProcessGroupNCCL)."""

    def __init__(self, handle, rank, nranks, device):
        self.handle = handle
        self.rank = rank
        self.nranks = nranks
        self.device = device
        self.comm_stream = core.Stream(handle=lib.prb_comm_stream(handle))

    @property
    def world_size(self):
        return self.nranks

    def all_reduce(self, tensor: Tensor, op: str = "sum", stream=None) -> Tensor:
        p = tensor.prb
        call("prb_comm_all_reduce", self.handle, C.byref(p), C.byref(p), {"sum": 0, "max": 1}[op],
             core._stream_handle(stream))
        return tensor

    def all_gather(self, tensor: Tensor, out: Tensor | None = None, stream=None) -> Tensor:
        """``out`` holds nranks x tensor.numel elements, rank r's contribution at r * tensor.numel."""
        if out is None:
            out = Tensor(tensor.numel * self.nranks, tensor.dtype, tensor.device)
        a, b = tensor.prb, out.prb
        call("prb_comm_all_gather", self.handle, C.byref(a), C.byref(b), core._stream_handle(stream))
        return out

    def broadcast(self, tensor: Tensor, src: int = 0, stream=None) -> Tensor:
        p = tensor.prb
        call("prb_comm_broadcast", self.handle, C.byref(p), int(src), core._stream_handle(stream))
        return tensor

    def barrier(self):
        t = Tensor.zeros(1, "float32", self.device)
        self.all_reduce(t)
        core.current_stream().synchronize()

    def __del__(self):
        try:
            lib.prb_comm_destroy(self.handle)
        except Exception:
            pass


_DONE = object()  # a sparse group whose exchange has run in this backward
_default_group: Group | None = None
_init_calls = 0  # every rank makes the same sequence of calls, so the count names the communicator


def init_parallel_env(store=None) -> Group:
    """Create the default communicator.  ref: parallel.py:921-1211, process_group_nccl.cc:559-571.

    Rank 0 creates the NCCL unique id and publishes it under the reference's key; peers fetch it.
    ``store`` may be any object with ``set(key, bytes)`` / ``get(key) -> bytes``.
    """
    global _default_group, _init_calls
    gid = _init_calls
    _init_calls += 1
    env = ParallelEnv()
    core.set_device(env.device_id)
    h = C.c_void_p()
    if env.world_size == 1:
        call("prb_comm_init_single", env.device_id, C.byref(h))
    else:
        if store is None:
            if env.under_torchrun:
                # the launcher owns MASTER_PORT; all workers share the agent as parent process
                path = os.path.join(tempfile.gettempdir(), "prb_store_%d_%d" % (os.getppid(), env.master_port))
                store = FileStore(path)
            else:
                store = TCPStore(env.master_addr, env.master_port, env.rank == 0, env.world_size)
        key = "ProcessGroupNCCL/nccl_ids/%d/0" % gid
        uid = (C.c_uint8 * _lib.UNIQUE_ID_BYTES)()
        if env.rank == 0:
            call("prb_comm_get_unique_id", uid)
            store.set(key, bytes(uid))
        else:
            raw = store.get(key)
            C.memmove(uid, raw, _lib.UNIQUE_ID_BYTES)
        call("prb_comm_init_rank", uid, env.world_size, env.rank, env.device_id, C.byref(h))
    _default_group = Group(h.value, env.rank, env.world_size, env.device_id)
    return _default_group


def get_group() -> Group | None:
    return _default_group


def get_rank() -> int:
    return _default_group.rank if _default_group else ParallelEnv().rank


def get_world_size() -> int:
    return _default_group.nranks if _default_group else ParallelEnv().world_size


def assign_group_by_size(tensors, is_sparse_gradient, group_size_limits=(25 * 1024 * 1024,), tensor_indices=()):
    """``core.eager_assign_group_by_size`` (ref distributed_py.cc:64-77, reducer.cc:45-156).

    ``tensors``: objects with ``numel`` and ``dtype`` (PRB dtype code).  Pure host logic: runs
    without a GPU.
    """
    n = len(tensors)
    if len(is_sparse_gradient) != n:  # ref reducer.cc:50-57
        raise _lib.PRBError(2, "tensors len must be equal to is_sparse_gradient len, but [%d] != [%d]"
                            % (n, len(is_sparse_gradient)))
    numels = (C.c_int64 * max(n, 1))(*[int(t.numel) for t in tensors])
    dtypes = (C.c_int32 * max(n, 1))(*[int(core.dtype_code(t.dtype)) for t in tensors])
    sparse = (C.c_uint8 * max(n, 1))(*[1 if s else 0 for s in is_sparse_gradient])
    limits = (C.c_uint64 * len(group_size_limits))(*[int(x) for x in group_size_limits])
    ti = (C.c_int64 * max(len(tensor_indices), 1))(*[int(x) for x in tensor_indices])
    flat = (C.c_int64 * max(n, 1))()
    sizes = (C.c_int64 * max(n, 1))()
    ng = C.c_int64()
    call("prb_assign_group_by_size", numels, dtypes, sparse, n, limits, len(group_size_limits), ti,
         len(tensor_indices), flat, sizes, C.byref(ng))
    out, k = [], 0
    for g in range(ng.value):
        out.append([int(flat[k + j]) for j in range(sizes[g])])
        k += sizes[g]
    return out


class EagerReducer:
    """``core.EagerReducer(tensors, group_indices, is_sparse, process_group, limits, find_unused)``.

    ref: reducer.h:88-142, distributed_py.cc:1285-1315.  ``process_group=None`` builds the
    host-only state machine (no device work), which is what CPU tests drive.
    """

    def __init__(self, grads, group_indices, is_sparse_gradient, process_group, group_size_limits=None,
                 find_unused_parameters=False):
        self._grads = list(grads)
        self.group_indices = [list(g) for g in group_indices]
        self._group = process_group
        # Sparse (SelectedRows) gradients: a sparse variable is always a group of its own (reducer.cc:94-99,543-551).
        # Those groups are exchanged from here (_allreduce_sparse); the native reducer sees the dense variables only,
        # renumbered 0..D-1, and its groups keep their relative order.  With no sparse variable nothing is renumbered.
        is_sparse = [bool(s) for s in is_sparse_gradient] if len(is_sparse_gradient) else [False] * len(self._grads)
        self._sparse_vars = [i for i, s in enumerate(is_sparse) if s]
        for g in self.group_indices:
            if any(is_sparse[v] for v in g) and len(g) != 1:
                raise _lib.PRBError(2, "a sparse gradient must be a group of its own, got a group of %d" % len(g))
        self._dense_of = {}
        for i in range(len(self._grads)):
            if not is_sparse[i]:
                self._dense_of[i] = len(self._dense_of)
        self._sparse_source = {v: (lambda v=v: self._grads[v]) for v in self._sparse_vars}
        self._sparse_ready = {}
        dense_groups = [[self._dense_of[v] for v in g] for g in self.group_indices if not is_sparse[g[0]]]
        dense_grads = [self._grads[i] for i in sorted(self._dense_of)]
        self.handle = None
        if dense_groups or not self._sparse_vars:
            flat = [v for g in dense_groups for v in g]
            sizes = [len(g) for g in dense_groups]
            self._flat = (C.c_int64 * max(len(flat), 1))(*flat)
            self._sizes = (C.c_int64 * max(len(sizes), 1))(*sizes)
            h = C.c_void_p()
            call("prb_reducer_create", tensor_array(dense_grads), len(dense_grads), self._flat, self._sizes, len(sizes),
                 process_group.handle if process_group is not None else None, int(bool(find_unused_parameters)),
                 C.byref(h))
            self.handle = h.value
        self.dense_group_count = len(dense_groups)

    # ---- sparse groups
    def set_sparse_source(self, var_index: int, getter):
        """``getter()`` returns the variable's current SelectedRows gradient (a new object after every backward)."""
        self._sparse_source[var_index] = getter

    def _mark_sparse_ready(self, var_index: int):
        if self._sparse_vars and self.finalized:  # after finalize hooks are no-ops (reducer.cc:740-758)
            return
        sr = self._sparse_source[var_index]()
        # ref reducer.cc:858-888
        if sr is None:
            raise _lib.PRBError(3, "The sparse parameter[%d] should have gradient. Currently, DataParallel does not "
                                "support sparse parameters without generating gradients during training." % var_index)
        if not isinstance(sr, core.SelectedRows):
            raise _lib.PRBError(3, "The sparse parameter[%d] must have a selectedrows gradient." % var_index)
        if var_index in self._sparse_ready:
            raise _lib.PRBError(3, "The sparse parameter[%d] was marked ready twice in one backward" % var_index)
        self._sparse_ready[var_index] = sr

    def _maybe_reduce_sparse(self, stream=None):
        """Sparse groups are exchanged once every gradient of the backward is there, in variable order (the same
        order on every rank, whatever order the hooks fired in); the reference launches them in group order between
        the dense groups (reducer.cc:922-929).  Same result: groups are independent."""
        if len(self._sparse_ready) != len(self._sparse_vars) or not self._sparse_vars:
            return
        if self.handle is not None and self.dense_group_count and not lib.prb_reducer_is_finalized(self.handle):
            return
        for v in self._sparse_vars:
            if self._sparse_ready[v] is not _DONE:
                self._allreduce_sparse(self._sparse_ready[v], stream)
                self._sparse_ready[v] = _DONE

    def _allreduce_sparse(self, sr, stream=None):
        """ref EagerReducer::AllReduceSparse reducer.cc:1072-1277: values scaled by 1/nranks, then every rank ends
        with all ranks' rows and values concatenated in rank order (duplicates are merged later, by the optimizer).
        The reference gathers with all-gather when the ranks hold the same number of rows and with 2 x nranks
        broadcasts otherwise; here the ragged case pads to the longest rank and still uses one all-gather for the
        rows and one for the values.  Host-synchronous like the reference (the row indices live on the host)."""
        from .. import _C_ops
        g = self._group
        if g is None:  # host-only state machine (CPU tests): no device work
            return
        n = g.nranks
        if sr.rows.size:
            _C_ops.scale_(sr.value, 1.0 / n, 0.0, False, stream=stream)
        if n == 1:
            return
        st = core._stream_handle(stream)
        width = sr.value.numel // sr.rows.size if sr.rows.size else 0
        meta = Tensor.from_numpy(np.array([sr.rows.size, width], np.int64), device=g.device, stream=stream)
        counts = g.all_gather(meta, stream=stream).numpy(stream).reshape(n, 2)
        width = int(counts[:, 1].max())
        longest = int(counts[:, 0].max())
        if longest == 0:
            return
        rows = np.zeros(longest, np.int64)
        rows[:sr.rows.size] = sr.rows
        all_rows = g.all_gather(Tensor.from_numpy(rows, device=g.device, stream=stream), stream=stream).numpy(stream)
        if sr.rows.size == longest:
            mine = sr.value
        else:
            mine = Tensor.zeros(longest * width, sr.value.dtype, g.device, stream=stream)
            if sr.rows.size:
                call("prb_memcpy_d2d", mine.data_ptr, sr.value.data_ptr, sr.value.nbytes, st)
        gathered = g.all_gather(mine, stream=stream)
        if (counts[:, 0] == longest).all():
            value = gathered
        else:
            es = _lib.DTYPE_SIZE[sr.value.dtype]
            value = Tensor(int(counts[:, 0].sum()) * width, sr.value.dtype, g.device)
            at = 0
            for r in range(n):
                nb = int(counts[r, 0]) * width * es
                if nb:
                    call("prb_memcpy_d2d", value.data_ptr + at, gathered.data_ptr + r * longest * width * es, nb, st)
                at += nb
            call("prb_stream_synchronize", st)  # `gathered` and `mine` are released on return
        sr.rows = np.concatenate([all_rows[r * longest:r * longest + int(counts[r, 0])] for r in range(n)])
        sr.value = value

    def prepare_for_backward(self, unused_vars=()):
        self._sparse_ready = {}
        if self.handle is None:
            return
        call("prb_reducer_prepare_for_backward", self.handle)
        if len(unused_vars):
            if any(int(v) in self._sparse_source for v in unused_vars):
                raise _lib.PRBError(3, "a sparse parameter cannot be unused (reducer.cc:858-872)")
            arr = (C.c_int64 * len(unused_vars))(*[self._dense_of.get(int(v), int(v)) for v in unused_vars])
            call("prb_reducer_set_unused_vars", self.handle, arr, len(unused_vars))

    def mark_var_ready(self, var_index: int, stream=None):
        if var_index in self._sparse_source:
            self._mark_sparse_ready(int(var_index))
        else:
            call("prb_reducer_mark_var_ready", self.handle, self._dense_of.get(int(var_index), int(var_index)),
                 core._stream_handle(stream) if self._group is not None else None)
        if self._sparse_vars:
            self._maybe_reduce_sparse(stream)

    def mark_vars_ready(self, var_indices, stream=None):
        if self._sparse_vars:
            for v in var_indices:
                self.mark_var_ready(int(v), stream)
            return
        arr = var_indices if isinstance(var_indices, C.Array) else (C.c_int64 * len(var_indices))(*var_indices)
        call("prb_reducer_mark_vars_ready", self.handle, arr, len(arr),
             core._stream_handle(stream) if self._group is not None else None)

    def set_unpack(self, enabled: bool):
        if self.handle is None:  # sparse groups only
            return
        call("prb_reducer_set_unpack", self.handle, int(bool(enabled)))

    def set_grads_in_arena(self, enabled: bool):
        if self.handle is None:  # sparse groups only
            return
        call("prb_reducer_set_grads_in_arena", self.handle, int(bool(enabled)))

    def set_grad_epilogue(self, loss_scale=None, found_inf=None, partial_base=None, use_mp_type=False):
        """Producer-side pass 1 (see prb_reducer_set_grad_epilogue).  Returns (active, partials view); all arguments
        None switches it off."""
        if self.handle is None or self._sparse_vars:
            return False, None
        out = _lib.PRBTensor()
        active = C.c_int32()
        if loss_scale is None and partial_base is None:
            call("prb_reducer_set_grad_epilogue", self.handle, None, C.byref(out), C.byref(active))
            self._epilogue_keep = None
            return False, None
        e = _lib.GradEpilogue()
        keep = [e]
        if loss_scale is not None:
            ls, fi = loss_scale.prb, found_inf.prb
            e.loss_scale, e.found_inf = C.pointer(ls), C.pointer(fi)
            keep += [ls, fi]
        if partial_base is not None:
            pb = np.ascontiguousarray(partial_base, np.int32)
            assert pb.size == len(self._grads)
            e.want_norm, e.partial_base = 1, pb.ctypes.data
            keep.append(pb)
        e.use_mp_type = int(bool(use_mp_type))
        call("prb_reducer_set_grad_epilogue", self.handle, C.byref(e), C.byref(out), C.byref(active))
        self._epilogue_keep = keep
        return bool(active.value), Tensor(out.numel, out.dtype, out.device, data_ptr=out.data, owner=self)

    @property
    def allreduce_backend(self) -> str:
        return {0: "none", 1: "nccl", 2: "nvlink", 3: "nvls"}[lib.prb_reducer_allreduce_backend(self.handle)]

    def allreduce_group(self, group_index: int, stream=None):
        call("prb_reducer_allreduce_group", self.handle, int(group_index), core._stream_handle(stream))

    def comm_error(self) -> bool:
        e = C.c_int32()
        call("prb_reducer_comm_error", self.handle, C.byref(e))
        return bool(e.value)

    def set_sharded(self, enabled: bool) -> bool:
        """One reduce-scatter over all buckets instead of an allreduce per bucket (prb_reducer_set_sharded).
        Returns whether the reducer can do it (several ranks and the peer-memory exchange)."""
        if self.handle is None or self._sparse_vars:
            return False
        active = C.c_int32()
        call("prb_reducer_set_sharded", self.handle, int(bool(enabled)), C.byref(active))
        return bool(active.value)

    def shard_ranges(self, rank: int, nranks: int = 0):
        """[(variable, element offset, length)] owned by ``rank`` in a sharded step, in arena order (``nranks`` 0: the
        communicator's size; any other value answers for a hypothetical job of that size)."""
        cap = len(self._grads)
        var, off, ln = ((C.c_int64 * max(cap, 1))() for _ in range(3))
        n = C.c_int64()
        call("prb_reducer_shard_ranges", self.handle, int(rank), int(nranks), var, off, ln, cap, C.byref(n))
        return [(int(var[i]), int(off[i]), int(ln[i])) for i in range(n.value)]

    def set_shard_layout(self, per_bucket: bool):
        if self.handle is not None:
            call("prb_reducer_set_shard_layout", self.handle, int(bool(per_bucket)))

    def allgather_params(self, params, skip=None, stream=None):
        """Every rank pulls the ranges it does not own from their owners' arenas into ``params`` (one tensor per
        variable, prb_reducer_allgather_params)."""
        if self._group is not None and self._group.nranks > 1 and self.allreduce_backend != "nvlink":
            # NCCL / NVLS reducers: every owner broadcasts its ranges (a step that found an Inf broadcasts the
            # unchanged values)
            for r in range(self._group.nranks):
                for v, off, ln in self.shard_ranges(r):
                    self._group.broadcast(params[v].view(off, ln), r, stream)
            return
        key = tuple(p.data_ptr for p in params)
        if getattr(self, "_gather_key", None) != key:
            self._gather_arr, self._gather_key = tensor_array(params), key
        sk = skip.prb if skip is not None else None
        call("prb_reducer_allgather_params", self.handle, self._gather_arr, len(params),
             C.byref(sk) if sk is not None else None, core._stream_handle(stream))

    def set_deferred_wait(self, enabled: bool):
        if self.handle is None:  # sparse groups only
            return
        call("prb_reducer_set_deferred_wait", self.handle, int(bool(enabled)))

    def wait_group(self, group_index: int, stream=None):
        call("prb_reducer_wait_group", self.handle, int(group_index), core._stream_handle(stream))

    def group_of(self, var_index: int) -> int:
        """Launch-order index of the variable's DENSE group (what wait_group / allreduce_group take)."""
        g = C.c_int64()
        call("prb_reducer_group_of", self.handle, self._dense_of.get(int(var_index), int(var_index)), C.byref(g))
        return int(g.value)

    @property
    def finalized(self) -> bool:
        dense = bool(lib.prb_reducer_is_finalized(self.handle)) if self.handle is not None else True
        return dense and all(self._sparse_ready.get(v) is _DONE for v in self._sparse_vars)

    def launched_groups(self):
        cap = self.dense_group_count
        out = (C.c_int64 * max(cap, 1))()
        n = C.c_int64()
        call("prb_reducer_launched_groups", self.handle, out, cap, C.byref(n))
        return [int(out[i]) for i in range(n.value)]

    def bucket_view(self, var_index: int) -> Tensor:
        p = _lib.PRBTensor()
        call("prb_reducer_bucket_view", self.handle, self._dense_of.get(int(var_index), int(var_index)), C.byref(p))
        return Tensor(p.numel, p.dtype, p.device, data_ptr=p.data, owner=self)

    def __del__(self):
        try:
            if self.handle is not None:
                lib.prb_reducer_destroy(self.handle)
        except Exception:
            pass


def sync_params_buffers(parameters, group: Group, src_rank: int = 0, comm_buffer_size_mb: int = 128):
    """Broadcast every parameter from ``src_rank`` at start-up, coalesced: same-dtype parameters are concatenated
    into ~128 MB buffers, each broadcast once and split back (ref parallel.py:156-199: build_groups with
    128 * 1024 * 1024, _coalesce_tensors, broadcast, _split_tensors) -- 3 collectives for GPT-2 medium instead of 292."""
    if group is None or group.nranks == 1:
        return
    from .. import _C_ops
    from .fleet_utils import build_groups
    params = [p for p in parameters if p.numel > 0]
    if not params:
        return
    for grp in build_groups(params, comm_buffer_size_mb * 1024 * 1024):
        if len(grp) == 1:
            group.broadcast(grp[0], src_rank)
            continue
        flat = _C_ops.concat(grp)
        group.broadcast(flat, src_rank)
        _C_ops.split_(flat, grp)
    core.current_stream().synchronize()


class DataParallel:
    """``paddle.DataParallel`` over a parameter list.

    Same constructor knobs and defaults as the reference (parallel.py:363-371): 25 MB buckets, a
    1 MB bucket for the parameters whose gradients arrive last, group order reversed so that the
    first gradients produced by backward fill group 0 (:478-491).

    Per step::

        dp.prepare_for_backward()          # what DataParallel.forward does (:548-561)
        ... for each gradient as it is produced:  dp.mark_var_ready(i)
        # when the last group has been launched the reducer finalises on its own

    ``consume_from_buckets=True`` keeps the reduced gradients in the bucket arena (no split back);
    ``dp.bucket_grads()`` then yields the views the optimizer reads directly.
    """

    def __init__(self, parameters, strategy=None, comm_buffer_size=25, last_comm_buffer_size=1,
                 find_unused_parameters=False, group=None, consume_from_buckets=False, grads_in_arena=False):
        self._parameters = [p for p in parameters if not getattr(p, "stop_gradient", False)]
        assert len(self._parameters) > 0, (
            "This model does not have any parameters to train, and does not need to use DataParallel")
        self.find_unused_parameters = find_unused_parameters
        self.grad_need_sync = True
        self.group = group if group is not None else _default_group
        if self.group is None:
            warnings.warn("init_parallel_env() was not called: falling back to a single-rank communicator.")
            self.group = init_parallel_env()
        self.grads_in_arena = bool(grads_in_arena)
        for p in self._parameters:
            if p.grad is None and not self.grads_in_arena and not getattr(p, "is_sparse_grad", False):
                p.grad = Tensor.zeros(p.numel, p.dtype, p.device)
        sync_params_buffers(self._parameters, self.group)
        self.comm_buffer_size = int(comm_buffer_size * 1024 * 1024)
        self.last_comm_buffer_size = int(last_comm_buffer_size * 1024 * 1024)
        self.init_reducer()
        self._consume = False
        self.pipelined = False
        self._epilogue_key = None     # what the reducer's grad epilogue is currently configured for
        self._epilogue_armed = False  # True once a backward has run with it (its results are there to consume)
        self._epilogue_partials = None
        self._sharded = False         # the exchange is a reduce-scatter from the NEXT backward on (set_sharded)
        self._sharded_armed = False   # ... and the backward that just ended ran that way
        if consume_from_buckets or self.grads_in_arena:
            self.consume_from_buckets(True)

    def init_reducer(self):
        is_sparse = [bool(getattr(p, "is_sparse_grad", False)) for p in self._parameters]
        self.group_indices = assign_group_by_size(self._parameters, is_sparse,
                                                  [self.last_comm_buffer_size, self.comm_buffer_size])
        if self.grads_in_arena:
            # gradients are PRODUCED in the bucket slots (no pack, no split back): p.grad becomes the slot view.
            # ref: MarkVarReady aliases the gradient into the group's dense tensor, reducer.cc:774-901
            grads = [Tensor(p.numel, p.dtype, p.device, data_ptr=0) for p in self._parameters]
        else:
            grads = [p.grad for p in self._parameters]
        self._reducer = EagerReducer(grads, list(reversed(self.group_indices)), is_sparse,
                                     self.group, [self.last_comm_buffer_size, self.comm_buffer_size],
                                     self.find_unused_parameters)
        self._sparse = is_sparse
        for i, p in enumerate(self._parameters):
            if is_sparse[i]:  # the producer hangs a new SelectedRows on p.grad in every backward
                self._reducer.set_sparse_source(i, lambda p=p: p.grad)
            elif self.grads_in_arena:
                p.grad = self._reducer.bucket_view(i)
        # backward produces gradients in reverse parameter order
        self._reverse_order = (C.c_int64 * len(self._parameters))(*reversed(range(len(self._parameters))))

    def consume_from_buckets(self, enabled: bool = True):
        """Leave the reduced gradients in the bucket arena and let the optimizer read them there.

        Each parameter gets ``_reduced_grad`` (the view of its slot) and ``_bucket`` (this wrapper and the
        launch-order index of its group).  With more than one rank the calc stream no longer waits for every
        bucket at the end of backward; the optimizer waits per bucket instead (``wait_group``)."""
        self._consume = bool(enabled)
        self._reducer.set_unpack(not self._consume)
        self.pipelined = self._consume and self.group.nranks > 1
        self._reducer.set_deferred_wait(self.pipelined)
        self._bucket_grads = [None if self._sparse[i] else self._reducer.bucket_view(i)
                              for i in range(len(self._parameters))] if enabled else None
        for i, p in enumerate(self._parameters):
            if self._sparse[i]:
                continue
            if enabled:
                p._reduced_grad = self._bucket_grads[i]
                p._bucket = (self, self._reducer.group_of(i))
            else:
                p._reduced_grad = None
                p._bucket = (None, 0)

    def set_sharded(self, enabled: bool = True) -> bool:
        """From the next backward on the buckets are reduce-scattered: each rank ends up with the reduced gradient
        of the arena ranges it owns only (``shard_ranges``) -- the exchange half of a sharded optimizer step
        (``GroupShardedOptimizerStage2``).  Needs ``consume_from_buckets``; returns False when the reducer cannot do
        it (one rank, NCCL / NVLS backend, sparse parameters), in which case nothing changes."""
        if enabled and not self._consume:
            raise _lib.PRBError(3, "set_sharded: the optimizer must read the buckets (consume_from_buckets=True)")
        self._sharded = self._reducer.set_sharded(enabled)
        return self._sharded

    def set_shard_layout(self, per_bucket: bool):
        """False (default): a rank owns one n-th of the whole arena, one reduce-scatter after the last bucket.  True: a
        rank owns one n-th of every bucket, one reduce-scatter per bucket as soon as it is packed (they hide under
        the rest of the backward pass).  To be chosen before the sharded optimizer is built."""
        self._reducer.set_shard_layout(per_bucket)

    def shard_ranges(self, rank=None):
        return self._reducer.shard_ranges(self.group.rank if rank is None else rank)

    def allgather_params(self, skip=None, stream=None):
        self._reducer.allgather_params(self._parameters, skip, stream)

    def wait_group(self, group_index: int, stream=None):
        self._reducer.wait_group(group_index, stream)

    def wait_all_groups(self, stream=None):
        for gi in range(self._reducer.dense_group_count):
            self._reducer.wait_group(gi, stream)

    def bucket_grads(self):
        return self._bucket_grads

    def parameters(self):
        return self._parameters

    def prepare_for_backward(self, unused_vars=()):
        if self.grad_need_sync:
            self._reducer.prepare_for_backward(unused_vars)
            self._epilogue_armed = self._epilogue_key is not None
            self._sharded_armed = self._sharded

    def _configure_epilogue(self, key, loss_scale=None, found_inf=None, partial_base=None, use_mp_type=False):
        """Called by the optimizer between two backward passes.  key None switches the epilogue off.  Returns True when
        the reducer will carry pass 1 from the NEXT backward on."""
        if key == self._epilogue_key:
            return key is not None
        self._epilogue_armed = False
        if key is None:
            self._reducer.set_grad_epilogue()
            self._epilogue_key, self._epilogue_partials = None, None
            return False
        active, partials = self._reducer.set_grad_epilogue(loss_scale, found_inf, partial_base, use_mp_type)
        if not active:
            self._reducer.set_grad_epilogue()
            self._epilogue_key, self._epilogue_partials = None, None
            return False
        self._epilogue_key, self._epilogue_partials = key, partials
        return True

    def mark_var_ready(self, index: int, stream=None):
        if self.grad_need_sync:
            self._reducer.mark_var_ready(index, stream)

    def mark_all_ready(self, stream=None):
        """Every gradient is ready (a finished backward): hooks fire in reverse parameter order."""
        if self.grad_need_sync:
            self._reducer.mark_vars_ready(self._reverse_order, stream)

    class _NoSync:
        def __init__(self, dp):
            self.dp = dp

        def __enter__(self):
            self.prev = self.dp.grad_need_sync
            self.dp.grad_need_sync = False

        def __exit__(self, *exc):
            self.dp.grad_need_sync = self.prev

    def no_sync(self):
        """Gradient accumulation without communication (ref parallel.py:496-546)."""
        return DataParallel._NoSync(self)

    @property
    def finalized(self) -> bool:
        return self._reducer.finalized
