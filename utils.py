"""``from utils import set_seed, init_dist_env`` (reference utils.py:12-24)."""
from distributed_pytorch_from_scratch_b200.utils.dist_env import init_dist_env, pick_backend, set_seed  # noqa: F401
