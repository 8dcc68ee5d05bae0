"""paddle_ray_b200 -- B200-native data-parallel training-step tail behind Paddle's operator/API shapes.

Bucketed gradient sync (EagerReducer), AMP check_finite_and_unscale, squared_l2_norm global-norm
clip and adamw / adam / merged_adam / fused_adam, as hand-written sm_100a CUDA behind a C ABI
(include/paddle_ray_b200.h).  No PyTorch, no Triton, no CPU fallback.
"""
from . import _lib  # noqa: F401  (raises ImportError if the shared library is not built)
from . import _C_ops, core  # noqa: F401
from .core import Event, PinnedBuffer, SelectedRows, Stream, Tensor, current_stream, device_count, set_device, synchronize  # noqa: F401

from . import amp, distributed, nn, optimizer  # noqa: F401,E402
from .distributed import DataParallel  # noqa: F401,E402

__version__ = "0.1"
