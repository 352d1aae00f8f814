from .policy import ACTPolicy, DETRVAE  # noqa: F401
from .rollout import RolloutMtAct  # noqa: F401
