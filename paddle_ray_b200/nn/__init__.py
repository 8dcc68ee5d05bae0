from .clip import ClipGradByGlobalNorm  # noqa: F401
