from .policy import ACTPolicy, DETRVAE  # noqa: F401
from .rollout import RolloutAct  # noqa: F401
