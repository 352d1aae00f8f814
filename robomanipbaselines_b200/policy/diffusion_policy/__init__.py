from .policy import DiffusionUnetHybridImagePolicy  # noqa: F401
from .rollout import RolloutDiffusionPolicy  # noqa: F401
