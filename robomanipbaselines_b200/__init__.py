"""B200-native batched policy-rollout engine for RoboManipBaselines' ACT hot path."""

__version__ = "0.1.0"
