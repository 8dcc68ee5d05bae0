"""``paddle.nn.ClipGradByGlobalNorm``.  ref: python/paddle/nn/clip.py:577-736."""
from __future__ import annotations

from .. import _C_ops
from .._lib import FLOAT32
from ..core import SelectedRows, Tensor

_clip_by_global_norm_using_mp_type_flag = False


def _clip_by_global_norm_using_mp_type(*args):
    """ref clip.py:179-192: get / set the flag that casts 16-bit gradients to fp32 for the norm."""
    global _clip_by_global_norm_using_mp_type_flag
    assert len(args) <= 1
    if len(args) == 1:
        assert isinstance(args[0], bool)
        old = _clip_by_global_norm_using_mp_type_flag
        _clip_by_global_norm_using_mp_type_flag = args[0]
        return old
    return _clip_by_global_norm_using_mp_type_flag


class ClipGradBase:
    def __call__(self, params_grads):
        return self._dygraph_clip(params_grads)


class ClipGradByGlobalNorm(ClipGradBase):
    """t_list[i] = t_list[i] * clip_norm / max(global_norm, clip_norm)   (ref clip.py:577-642).

    When handed to an optimizer of this package the clip runs fused inside the optimizer's two-pass
    step (``prb_dp_step_tail``).  Calling it directly reproduces ``_dygraph_clip``: new gradient
    tensors are returned, the originals are left untouched (ref :715-732).
    """

    def __init__(self, clip_norm, group_name="default_group", auto_skip_clip=False):
        self.clip_norm = float(clip_norm)
        self.group_name = group_name
        assert isinstance(auto_skip_clip, bool)
        self.auto_skip_clip = auto_skip_clip
        self.last_global_norm: Tensor | None = None
        self.last_clip: Tensor | None = None

    def _ensure_outputs(self):
        """The two device scalars every user of this object writes into; allocated once, never rebound (cached
        optimizer plans hold their addresses)."""
        if self.last_global_norm is None:
            self.last_global_norm, self.last_clip = Tensor.zeros(1, FLOAT32), Tensor.zeros(1, FLOAT32)

    def __str__(self):
        return "Gradient Clip By GlobalNorm, global_norm=%f" % (self.clip_norm)

    def _dygraph_clip(self, params_grads):
        todo = [(p, g) for p, g in params_grads if g is not None and getattr(p, "need_clip", True) is not False]
        if not todo:
            return params_grads  # "all parameters have been filtered out" (ref :676-682)
        new_grads, work, sparse_values = {}, [], []
        for p, g in todo:
            if isinstance(g, SelectedRows):
                # the norm of a sparse gradient is taken from its MERGED rows (ref :655-662); the gradient itself is
                # scaled as it is, duplicates and all (paddle.multiply(g, clip_input), :715-727)
                work.append(_C_ops.merge_selected_rows(g).value)
                nv = Tensor(max(g.value.numel, 1), g.value.dtype, g.value.device).view(0, g.value.numel)
                if g.value.numel:
                    nv.copy_(g.value)
                sparse_values.append(nv)
                new_grads[id(p)] = SelectedRows(g.rows.copy(), g.height, nv)
            else:
                ng = Tensor(g.numel, g.dtype, g.device)
                ng.copy_(g)
                new_grads[id(p)] = ng
                work.append(ng)
        self._ensure_outputs()
        _C_ops.clip_grad_by_global_norm_(work, self.clip_norm, self.auto_skip_clip,
                                         _clip_by_global_norm_using_mp_type(), norm=self.last_global_norm,
                                         clip=self.last_clip)
        sparse_values = [v for v in sparse_values if v.numel]
        if sparse_values:
            _C_ops.multiply_by_clip_(sparse_values, self.last_clip)
        return [(p, new_grads.get(id(p), g)) for p, g in params_grads]
