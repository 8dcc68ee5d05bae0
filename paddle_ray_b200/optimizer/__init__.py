from . import lr  # noqa: F401
from .adam import Adam, AdamW  # noqa: F401
from .optimizer import Optimizer  # noqa: F401
