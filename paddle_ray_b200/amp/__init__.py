from .grad_scaler import AmpScaler, GradScaler, OptimizerState  # noqa: F401
