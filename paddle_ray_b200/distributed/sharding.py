"""Sharded optimizer state (ZeRO stage 1) -- ``DygraphShardingOptimizer``.

ref: python/paddle/distributed/fleet/meta_optimizers/dygraph_optimizer/dygraph_sharding_optimizer.py:28-213
     (greedy partition of whole parameters over the sharding ranks :101-126, inner optimizer over the local
     shard :141-149, parameter broadcast from the owner after every step :151-171,192-200).

Every rank still receives the full reduced gradients (the DataParallel bucket exchange); only the rank that owns a
parameter keeps its moments / master weight and runs its update, so the optimizer's state and HBM traffic per rank
drop from 28 B/elem to 28/N, and the updated parameters travel once over NVLink.  B200-first difference from the
reference: the owner's parameters are synchronised COALESCED -- same-dtype parameters of one owner are concatenated,
broadcast with one collective and split back (the reference issues one broadcast per parameter, :158-171).
Because each parameter is updated by exactly one rank with the same kernel the replicated optimizer would run, the
result is bit-identical to the replicated run on every rank (tests/dist_worker.py).

Not supported here: a global-norm gradient clip (the reference routes it through HybridParallelClipGrad, which
allreduces the per-shard squared norms; outside this path).
"""
from __future__ import annotations

from functools import reduce

import numpy as np

from .. import _C_ops, _lib, core
from . import parallel as _parallel
from .fleet_utils import build_groups


class ShardingGroup:
    """The three questions DygraphShardingOptimizer asks its ``hcg`` (HybridCommunicateGroup), answered from a
    communicator of this package."""

    def __init__(self, group=None):
        self._group = group if group is not None else _parallel.get_group()

    def get_sharding_parallel_world_size(self):
        return self._group.nranks if self._group is not None else 1

    def get_sharding_parallel_rank(self):
        return self._group.rank if self._group is not None else 0

    def get_sharding_parallel_group(self):
        return self._group


class DygraphShardingOptimizer:
    def __init__(self, hcg, user_defined_strategy, params, inner_optimizer_class, **inner_optimizer_kargs):
        if not isinstance(params, list):
            raise TypeError("`parameters` argument given to the DygraphShardingOptimizer should be an iterable of "
                            "paddle Tensors, but got argument type is `{}`.".format(type(params)))
        if inner_optimizer_kargs.get("grad_clip") is not None:
            raise _lib.PRBError(4, "DygraphShardingOptimizer: a global-norm clip over sharded parameters needs the "
                                   "hybrid-parallel clip (allreduce of per-shard norms), which is outside this path")
        self._parameter_list = params
        for i, p in enumerate(params):
            if p.name is None:
                p.name = "param_%d" % i
        self._inner_optimizer_class = inner_optimizer_class
        self._inner_optimizer_kargs = inner_optimizer_kargs
        self._hcg = hcg if hcg is not None else ShardingGroup()
        self._user_defined_strategy = user_defined_strategy
        self._sharding_world_size = self._hcg.get_sharding_parallel_world_size()
        self._sharding_rank = self._hcg.get_sharding_parallel_rank()
        self._build_sharding_mapping()
        self._buid_inner_optimizer()
        self._sync_plan = None

    def clear_grad(self):
        for p in self._parameter_list:
            if not p.stop_gradient:
                p.clear_gradient()

    def _build_sharding_mapping(self):
        self._rank2params = self._partition_parameters()
        self._param2rank = self._map_param_to_rank()

    def _partition_parameters(self):
        """Greedy: each parameter goes to the rank that holds the fewest elements so far (ref :101-126)."""
        mapping = {r: [] for r in range(self._sharding_world_size)}
        sizes = [0] * self._sharding_world_size
        for param in self._parameter_list:
            rank = sizes.index(min(sizes))
            mapping[rank].append(param)
            numel = reduce(lambda x, y: x * y, param.shape)
            assert numel > 0, "param [{}] should larger than 0, but it is [{}]".format(param.name, numel)
            sizes[rank] += numel
        return mapping

    def _map_param_to_rank(self):
        return {p.name: rank for rank, params in self._rank2params.items() for p in params}

    def _buid_inner_optimizer(self):
        self._inner_optimizer = self._inner_optimizer_class(parameters=self._rank2params[self._sharding_rank],
                                                            **self._inner_optimizer_kargs)

    def _sharding_sync_parameters(self):
        """Every owner broadcasts its updated parameters (ref :151-171), coalesced per owner and dtype."""
        group = self._hcg.get_sharding_parallel_group()
        if group is None or self._sharding_world_size == 1:
            return
        if self._sync_plan is None:
            self._sync_plan = []
            for rank, params in self._rank2params.items():
                if not params:
                    continue
                for grp in build_groups(params, 128 * 1024 * 1024):
                    flat = None if len(grp) == 1 else core.Tensor(sum(p.numel for p in grp), grp[0].dtype, grp[0].device)
                    self._sync_plan.append((rank, grp, flat))
        for rank, grp, flat in self._sync_plan:
            if flat is None:
                group.broadcast(grp[0], rank)
                continue
            if rank == self._sharding_rank:
                o = flat.prb
                _lib.call("prb_concat", _lib.tensor_array(grp), len(grp), _lib.C.byref(o), core._stream_handle(None))
            group.broadcast(flat, rank)
            if rank != self._sharding_rank:
                _C_ops.split_(flat, grp)

    def minimize(self, loss=None, startup_program=None, parameters=None, no_grad_set=None):
        self.step()

    def step(self):
        self._inner_optimizer.step()
        self._sharding_sync_parameters()

    @property
    def _grad_clip(self):
        return self._inner_optimizer._grad_clip

    def __getattr__(self, item):
        if item == "_inner_optimizer":
            raise AttributeError(item)
        return getattr(self._inner_optimizer, item)


class GroupShardedOptimizerStage2:
    """Optimizer state AND update sharded over the data-parallel ranks, on top of the reducer's own exchange.

    ref: python/paddle/distributed/fleet/meta_parallel/sharding/group_sharded_optimizer_stage2.py:54-612 with
         group_sharded_stage2.py (gradients reduced to the rank that owns them, owner-only update, parameters
         broadcast from the owner).  Same constructor shape: ``GroupShardedOptimizerStage2(params, optim, group)``.

    The reference shards by PARAMETER (greedy by size) and moves data with one ``reduce`` per gradient plus one
    ``broadcast`` per parameter.  Here ownership is positional: rank r owns the r-th n-th of the bucket arena's
    8192-element chunks, whatever tensors they belong to, so the shards are equal to within one chunk and the whole
    step is

        pack (all buckets, one launch) -> ONE reduce-scatter kernel over NVLink peer memory, which also leaves the
        inf/nan flag and the squared-norm partials of the reduced gradient on every rank -> norm finalize -> update
        of the owned ranges only (moments and master weights exist for those ranges only: 12 B/elem / nranks), the new
        parameter values written over their gradients in the arena -> ONE gather kernel in which every rank pulls
        the other ranks' ranges straight into its parameters.

    NVLink traffic equals the allreduce's (a two-shot allreduce IS these two halves); the update pass shrinks from
    28 B/elem to 28/nranks.  Every element is updated by exactly one rank with the kernel and the clip factor the
    replicated optimizer would use, so parameters are bit-identical to the replicated run on every rank
    (tests/dist_worker.py).

    The FIRST step after construction still exchanges with the allreduce (the reducer's check + norm epilogue and the
    reduce-scatter are switched on at its end, between two backward passes); its update is already sharded.
    The reduced gradients are consumed: after a step the arena holds parameter values, ``p.grad`` is not written.
    """

    def __init__(self, params, optim, group=None, offload=False, device="gpu", pertrain_sync_models=True,
                 dp_group=None, reduce_overlap=False, **kw):
        """``reduce_overlap`` (ref ``_set_reduce_overlap``, group_sharded_optimizer_stage2.py:237-239): True shards every
        BUCKET over the ranks and reduce-scatters each bucket as soon as it is packed, so that all but the last
        exchange hide under the rest of the backward pass; False (default) shards the arena as a whole and runs ONE
        reduce-scatter after the last bucket (fewer launches and barriers when there is nothing to hide under)."""
        if offload:
            raise _lib.PRBError(4, "GroupShardedOptimizerStage2: offload is not built")
        self._all_params = list(params)
        self._optim = optim
        if optim._created:
            raise _lib.PRBError(3, "GroupShardedOptimizerStage2: wrap the optimizer before its first step")
        dp = self._all_params[0]._bucket[0]
        if dp is None or any(p._bucket[0] is not dp for p in self._all_params) or \
                len(self._all_params) != len(dp._parameters):
            raise _lib.PRBError(3, "GroupShardedOptimizerStage2: the parameters must be exactly those of ONE "
                                   "DataParallel(..., consume_from_buckets=True)")
        if len({p.dtype for p in self._all_params}) != 1:
            raise _lib.PRBError(4, "GroupShardedOptimizerStage2: parameters of one dtype only")
        self._dp = dp
        dp.set_shard_layout(bool(reduce_overlap))
        self._group = group if group is not None else dp.group
        self._rank, self._nranks = self._group.rank, self._group.nranks
        # the ranges this rank owns, as parameter-like views (what the wrapped optimizer now optimizes)
        self._local_params = []
        self._param2range = {}
        for v, off, ln in dp.shard_ranges(self._rank):
            parent = dp._parameters[v]
            view = parent.view(off, ln)
            view.name = parent.name
            view.need_clip, view.optimize_attr, view.stop_gradient = parent.need_clip, parent.optimize_attr, False
            view._reduced_grad = parent._reduced_grad.view(off, ln)
            view._bucket = (None, 0)
            view._parent = parent
            self._local_params.append(view)
            self._param2range[parent.name] = (off, ln)
        if not self._local_params:
            raise _lib.PRBError(3, "GroupShardedOptimizerStage2: rank %d owns nothing (fewer 8192-element chunks than "
                                   "ranks)" % self._rank)
        optim._parameter_list = self._local_params   # ref _update_opt_status / _segment_params :301-328
        self._auxiliary_vars = {}
        self._stage1 = None
        self._stage2 = None
        self._first = True

    # ---- what GradScaler.step / user code touch
    def _set_auxiliary_var(self, key, val):
        self._auxiliary_vars[key] = val
        self._optim._set_auxiliary_var(key, val)

    def _get_auxiliary_var(self, key):
        return self._auxiliary_vars.get(key)

    @property
    def local_params(self):
        return self._local_params

    @property
    def _parameter_list(self):
        return self._all_params

    @property
    def _grad_clip(self):
        return self._optim._grad_clip

    def clear_grad(self, set_to_zero=True):
        for p in self._all_params:
            p.clear_gradient(set_to_zero)

    def get_lr(self):
        return self._optim.get_lr()

    def set_lr(self, lr):
        self._optim.set_lr(lr)

    def state_dict(self):
        """The wrapped optimizer's keys; moments and master weights cover this rank's range of each parameter only
        (``shard_ranges``), like the reference's per-rank optimizer state."""
        return self._optim.state_dict()

    def set_state_dict(self, state_dict):
        self._optim.set_state_dict(state_dict)

    def minimize(self):
        raise RuntimeError("optimizer.minimize() not support now, please use optimizer.step()")  # ref :550-553

    # ---- step
    def _stage1_args(self, found_inf, loss_scale, partials):
        """Argument block of the check + norm + clip-factor call over ALL gradients."""
        from ..nn.clip import _clip_by_global_norm_using_mp_type
        clip = self._optim._grad_clip
        cfg = _lib.TailConfig()
        cfg.stage = 1
        cfg.use_amp = int(loss_scale is not None)
        cfg.use_clip = int(clip is not None)
        cfg.clip_norm = float(clip.clip_norm) if clip else 0.0
        cfg.auto_skip_clip = int(clip.auto_skip_clip) if clip else 0
        cfg.use_mp_type = int(_clip_by_global_norm_using_mp_type())
        if partials is not None:
            cfg.pass1_done, cfg.pass1_partials = 1, partials.data_ptr
        keep = {"cfg": cfg, "params": _lib.tensor_array(self._all_params),
                "grads": _lib.tensor_array([p._reduced_grad for p in self._all_params]),
                "ls": loss_scale.prb if loss_scale is not None else None,
                "fi": found_inf.prb if found_inf is not None else None, "gn": None, "cl": None}
        if clip is not None:
            clip._ensure_outputs()
            keep["gn"], keep["cl"] = clip.last_global_norm.prb, clip.last_clip.prb
        C = _lib.C
        ref = lambda k: C.byref(keep[k]) if keep[k] is not None else None  # noqa: E731
        return (C.byref(cfg), keep["params"], keep["grads"], None, None, None, None, None, None, len(self._all_params),
                None, None, ref("ls"), ref("fi"), ref("gn"), ref("cl"), keep)

    def step(self):
        from ..nn.clip import _clip_by_global_norm_using_mp_type
        opt, dp, C = self._optim, self._dp, _lib.C
        opt._create_accumulators()
        found_inf = self._get_auxiliary_var("found_inf")
        loss_scale = self._get_auxiliary_var("loss_scale")
        clip = opt._grad_clip
        if found_inf is not None and loss_scale is None:
            raise _lib.PRBError(4, "GroupShardedOptimizerStage2: call scaler.step(optimizer) without a separate "
                                   "scaler.unscale_() (the check has to see the reduced gradients of every rank)")
        needs_pass1 = clip is not None or loss_scale is not None
        mp = _clip_by_global_norm_using_mp_type()
        key = (id(self), loss_scale.data_ptr if loss_scale is not None else 0, clip is not None, mp)
        pass1_done = needs_pass1 and dp._epilogue_key == key and dp._epilogue_armed
        if dp._sharded_armed and needs_pass1 and not pass1_done:
            raise _lib.PRBError(3, "GroupShardedOptimizerStage2: the AMP / clip configuration changed after a sharded "
                                   "backward; the gradients of this step were only reduced on their owners")
        dp.wait_all_groups()
        stream = core._stream_handle(None)
        plan_key = (key, pass1_done, found_inf.data_ptr if found_inf is not None else 0, opt._hyper_key(),
                    (clip.clip_norm, clip.auto_skip_clip) if clip is not None else None)
        if self._stage2 is None or self._stage2[0] != plan_key:
            s1 = self._stage1_args(found_inf, loss_scale, dp._epilogue_partials if pass1_done else None) \
                if needs_pass1 else None
            pg = [(p, p._reduced_grad) for p in self._local_params]
            s2 = opt._tail_args(pg, found_inf, loss_scale, keep_beta_pow=False)
            cfg = s2[-1]["cfg"]
            cfg.stage, cfg.param_to_grad_slot, cfg.write_back_grads, cfg.zero_grads = 2, 1, 0, 0
            self._stage1, self._stage2 = s1, (plan_key, s2)
        if self._stage1 is not None:
            _lib.call("prb_dp_step_tail", *self._stage1[:16], stream)
        lr = opt._global_lr().prb
        a = self._stage2[1]
        _lib.call("prb_dp_step_tail", a[0], a[1], a[2], C.byref(lr), *a[4:16], stream)
        dp.allgather_params(found_inf)
        if self._first:
            # between two backward passes: from the next one on the exchange is the reduce-scatter and the reducer's
            # exchange kernel carries the check and the norm partials
            self._first = False
            if needs_pass1:
                chunk = _lib.lib.prb_grad_chunk_elems()
                base = np.zeros(len(self._all_params), np.int32)
                at = 0
                for i, p in enumerate(self._all_params):   # == dp._parameters order
                    base[i] = at
                    at += (p.numel + chunk - 1) // chunk
                if not dp._configure_epilogue(key, loss_scale, found_inf, base if clip is not None else None, mp):
                    return  # this reducer cannot carry the epilogue: stay with the allreduce + sharded update
            dp.set_sharded(True)
