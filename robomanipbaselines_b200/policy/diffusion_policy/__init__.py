from .policy import DiffusionUnetHybridImagePolicy  # noqa: F401
