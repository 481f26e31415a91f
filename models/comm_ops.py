"""``from models.comm_ops import Split, Reduce, Copy, Gather`` (reference models/comm_ops.py:7-83; implementation in
distributed_pytorch_from_scratch_b200/parallel/comm_ops.py)."""
from distributed_pytorch_from_scratch_b200.parallel.comm_ops import (  # noqa: F401
    Copy, Gather, Reduce, Split, all_gather_last_dim, all_reduce_sum, local_chunk_last_dim)
