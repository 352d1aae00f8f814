from .policy import DP3  # noqa: F401
