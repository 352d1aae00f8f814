from .policy import ACTPolicy, DETRVAE  # noqa: F401
