"""Seeding, distributed bootstrap, checkpoints, synthetic data and timers."""
from .dist_env import init_dist_env, set_seed  # noqa: F401
