from .selective_adam import SelectiveAdam  # noqa: F401
