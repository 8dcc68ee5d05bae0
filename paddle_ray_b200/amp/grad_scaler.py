"""``paddle.amp.GradScaler`` / ``AmpScaler`` with all state on the device.

ref: python/paddle/amp/grad_scaler.py:88-97,654-663 (defaults), :253-399 (_unscale), :401-428
     (_update), :543-600 (state_dict), :742-830 (step / update / minimize).

The reference branches on ``found_inf`` on the host (one D2H sync per step) and keeps the
loss-scale state machine in Python.  Here ``found_inf``, the loss scale and both counters are device
scalars; ``update()`` is the ``update_loss_scaling`` kernel (ref impl/amp_kernel_impl.h:38-75), so a
training step never synchronises.  ``step(optimizer)`` hands the scale to an optimizer of this
package, which unscales, checks and updates in one fused two-pass call.
"""
from __future__ import annotations

from collections import defaultdict
from enum import Enum

import numpy as np

import ctypes as C

from .. import _C_ops, core
from .._lib import BOOL, FLOAT32, INT32, call
from ..core import Tensor


class OptimizerState(Enum):
    INIT = 0
    UNSCALED = 1
    STEPPED = 2


def _refresh_optimizer_state():
    return {"state": OptimizerState.INIT}


class AmpScaler:
    def __init__(self, enable=True, init_loss_scaling=2.0 ** 15, incr_ratio=2.0, decr_ratio=0.5,
                 incr_every_n_steps=1000, decr_every_n_nan_or_inf=1, use_dynamic_loss_scaling=True):
        self._enable = enable
        if self._enable:
            assert incr_ratio > 1.0, "The incr_ratio must be > 1.0."
            assert decr_ratio < 1.0, "The decr_ratio must be < 1.0."
            self._init_loss_scaling = init_loss_scaling
            self._incr_ratio = incr_ratio
            self._decr_ratio = decr_ratio
            self._incr_every_n_steps = incr_every_n_steps
            self._decr_every_n_nan_or_inf = decr_every_n_nan_or_inf
            self._use_dynamic_loss_scaling = use_dynamic_loss_scaling
            self._found_inf = Tensor.zeros(1, BOOL)
            self._temp_found_inf = Tensor.zeros(1, BOOL)
            self._scale = Tensor.from_numpy(np.array([init_loss_scaling], np.float32))
            self._good = Tensor.zeros(1, INT32)  # incr_count
            self._bad = Tensor.zeros(1, INT32)   # decr_count
            self._optimizer_states = defaultdict(_refresh_optimizer_state)

    def scale(self, var: Tensor) -> Tensor:
        """``var * loss_scale`` (ref :170-200); a new tensor."""
        if not self._enable:
            return var
        out = Tensor(var.numel, var.dtype, var.device)
        a, s, o = var.prb, self._scale.prb, out.prb
        call("prb_multiply_by_scalar_tensor", C.byref(a), C.byref(s), C.byref(o), core._stream_handle(None))
        return out

    def get_loss_scaling(self) -> Tensor:
        return self._scale

    # ---- unscale (ref :253-399)
    def _unscale(self, optimizer):
        if not self._enable:
            return
        state = self._optimizer_states[id(optimizer)]
        if state["state"] is OptimizerState.UNSCALED:
            raise RuntimeError("unscale_() has already been called on this optimizer since the last update().")
        if state["state"] is OptimizerState.STEPPED:
            raise RuntimeError("unscale_() is being called after step().")
        by_dtype = {}
        for p in optimizer._parameter_list:
            g = getattr(p, "_reduced_grad", None) or p.grad
            if g is not None:
                by_dtype.setdefault(g.dtype, []).append(g)
        self._found_inf.zero_()
        for grads in by_dtype.values():
            _C_ops.check_finite_and_unscale_(grads, self._scale, self._temp_found_inf)
            from ..optimizer.optimizer import _or_into
            _or_into(self._found_inf, self._temp_found_inf)
        state["state"] = OptimizerState.UNSCALED

    def unscale_(self, optimizer):
        return self._unscale(optimizer)

    # ---- step / update (ref :742-830)
    def step(self, optimizer):
        if not self._enable:
            return optimizer.step()
        state = self._optimizer_states[id(optimizer)]
        if state["state"] is OptimizerState.STEPPED:
            raise RuntimeError("step() has already been called since the last update().")
        optimizer._set_auxiliary_var("found_inf", self._found_inf)
        # INIT: the optimizer unscales + checks inside its fused step; UNSCALED: skip flag only
        optimizer._set_auxiliary_var("loss_scale", self._scale if state["state"] is OptimizerState.INIT else None)
        optimizer.step()
        optimizer._set_auxiliary_var("loss_scale", None)
        state["state"] = OptimizerState.STEPPED
        if not self._use_dynamic_loss_scaling:
            self._optimizer_states = defaultdict(_refresh_optimizer_state)

    def minimize(self, optimizer, *args, **kwargs):
        self.step(optimizer)
        if self._enable and self._use_dynamic_loss_scaling:
            self._update()
        self._optimizer_states = defaultdict(_refresh_optimizer_state)
        return None, None

    def _update(self):
        """ref :401-428, executed on the device (ref amp_kernel.cu:84-107)."""
        if not self._enable:
            return
        _C_ops.update_loss_scaling_([], self._found_inf, self._scale, self._good, self._bad, self._incr_every_n_steps,
                                    self._decr_every_n_nan_or_inf, self._incr_ratio, self._decr_ratio, False)

    def update(self):
        if not self._enable:
            return
        if self._use_dynamic_loss_scaling:
            self._update()
            self._optimizer_states = defaultdict(_refresh_optimizer_state)

    # ---- introspection (these read device state back: not for the hot loop)
    def is_enable(self):
        return self._enable

    def is_use_dynamic_loss_scaling(self):
        return self._use_dynamic_loss_scaling

    def get_init_loss_scaling(self):
        return self._init_loss_scaling

    def state_dict(self):
        """ref :543-573, same keys."""
        if not self._enable:
            return {}
        return {
            "scale": self._scale.numpy(), "incr_ratio": self._incr_ratio, "decr_ratio": self._decr_ratio,
            "incr_every_n_steps": self._incr_every_n_steps, "decr_every_n_nan_or_inf": self._decr_every_n_nan_or_inf,
            "incr_count": int(self._good.item()), "decr_count": int(self._bad.item()),
            "use_dynamic_loss_scaling": self._use_dynamic_loss_scaling,
        }

    def load_state_dict(self, state_dict):
        """ref :575-600."""
        if not self._enable:
            return
        if len(state_dict) == 0:
            raise RuntimeError("The input state dict is empty, possibly because it was saved from a disabled "
                               "instance of GradScaler.")
        self._init_loss_scaling = state_dict["scale"][0]
        self._scale.copy_from_numpy(np.array([self._init_loss_scaling], np.float32))
        self._incr_ratio = state_dict["incr_ratio"]
        self._decr_ratio = state_dict["decr_ratio"]
        self._incr_every_n_steps = state_dict["incr_every_n_steps"]
        self._decr_every_n_nan_or_inf = state_dict["decr_every_n_nan_or_inf"]
        self._good.copy_from_numpy(np.array([state_dict["incr_count"]], np.int32))
        self._bad.copy_from_numpy(np.array([state_dict["decr_count"]], np.int32))
        self._use_dynamic_loss_scaling = state_dict["use_dynamic_loss_scaling"]


class GradScaler(AmpScaler):
    """ref :654-663: same as AmpScaler with decr_every_n_nan_or_inf defaulting to 2."""

    def __init__(self, enable=True, init_loss_scaling=2.0 ** 15, incr_ratio=2.0, decr_ratio=0.5,
                 incr_every_n_steps=1000, decr_every_n_nan_or_inf=2, use_dynamic_loss_scaling=True):
        super().__init__(enable, init_loss_scaling, incr_ratio, decr_ratio, incr_every_n_steps,
                         decr_every_n_nan_or_inf, use_dynamic_loss_scaling)
