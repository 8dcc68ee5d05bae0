"""``paddle.optimizer.Adam``.  ref: python/paddle/optimizer/adam.py:163-224 (ctor), :530-712 (multi-tensor)."""
from __future__ import annotations

import numpy as np

from .. import _C_ops
from .lr import LRScheduler  # noqa: F401
from .optimizer import Optimizer


class Adam(Optimizer):
    _use_adamw = False

    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8, parameters=None, weight_decay=None,
                 grad_clip=None, lazy_mode=False, multi_precision=False, use_multi_tensor=False, name=None):
        assert learning_rate is not None and beta1 is not None and beta2 is not None and epsilon is not None
        if not 0 <= beta1 < 1:
            raise ValueError("Invaild value of beta1, expect beta1 in [0,1).")
        if not 0 <= beta2 < 1:
            raise ValueError("Invaild value of beta2, expect beta2 in [0,1).")
        if not 0 <= epsilon:
            raise ValueError("Invaild value of epsilon, expect epsilon >= 0.")
        if weight_decay is not None:
            raise NotImplementedError("L2-regularisation via weight_decay is outside this path; use AdamW")
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name, multi_precision)
        self.type = "adam"
        self._beta1, self._beta2, self._epsilon = beta1, beta2, epsilon
        self._lazy_mode = lazy_mode
        self._use_multi_tensor = use_multi_tensor

    def _coeff(self):
        return 0.0

    def _append_optimize_op(self, p, g, lr, m1, m2, master, found_inf, lr_ratio, with_decay, use_global_beta_pow):
        """ref adam.py:390-470 -> _C_ops.adam_"""
        _C_ops.adam_(p, g, lr, m1, m2, self._beta1_pow, self._beta2_pow, master, found_inf, self._beta1, self._beta2,
                     self._epsilon, self._lazy_mode, 1000, master is not None, use_global_beta_pow)


class AdamW(Optimizer):
    """ref: python/paddle/optimizer/adamw.py:149-163 (signature and defaults)."""

    _use_adamw = True

    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8, parameters=None, weight_decay=0.01,
                 lr_ratio=None, apply_decay_param_fun=None, grad_clip=None, lazy_mode=False, multi_precision=False,
                 name=None):
        assert learning_rate is not None and beta1 is not None and beta2 is not None and epsilon is not None
        if not 0 <= beta1 < 1:
            raise ValueError("Invaild value of beta1, expect beta1 in [0,1).")
        if not 0 <= beta2 < 1:
            raise ValueError("Invaild value of beta2, expect beta2 in [0,1).")
        if not 0 <= epsilon:
            raise ValueError("Invaild value of epsilon, expect epsilon >= 0.")
        if not isinstance(weight_decay, float):
            raise TypeError("weight_decay should be float or Tensor.")
        if lr_ratio is not None:
            assert callable(lr_ratio)
        super().__init__(learning_rate, parameters, weight_decay, grad_clip, name, multi_precision)
        self.type = "adamw"
        self._beta1, self._beta2, self._epsilon = beta1, beta2, epsilon
        self._weight_decay = weight_decay
        self._lr_ratio = lr_ratio
        self._apply_decay_param_fun = apply_decay_param_fun
        self._lazy_mode = lazy_mode

    def _coeff(self):
        return self._weight_decay

    def _tensor_attrs(self, pg):
        """Per-parameter lr_ratio and with_decay (ref adamw.py:395-402,431-435)."""
        lr_ratio = np.ones(len(pg), np.float32)
        with_decay = np.ones(len(pg), np.uint8)
        for i, (p, _) in enumerate(pg):
            if self._lr_ratio is not None:
                lr_ratio[i] = self._lr_ratio(p)
            if self._apply_decay_param_fun is not None and not self._apply_decay_param_fun(p.name):
                with_decay[i] = 0
        return lr_ratio, with_decay

    def _append_optimize_op(self, p, g, lr, m1, m2, master, found_inf, lr_ratio, with_decay, use_global_beta_pow):
        """ref adamw.py:445-465 -> _C_ops.adamw_"""
        _C_ops.adamw_(p, g, lr, m1, m2, self._beta1_pow, self._beta2_pow, master, found_inf, self._beta1, self._beta2,
                      self._epsilon, lr_ratio, self._weight_decay, with_decay, self._lazy_mode, 1000, master is not None,
                      use_global_beta_pow)
