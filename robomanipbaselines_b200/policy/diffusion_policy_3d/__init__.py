from .policy import DP3  # noqa: F401
from .rollout import RolloutDiffusionPolicy3d  # noqa: F401
