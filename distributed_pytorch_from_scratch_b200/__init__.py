"""distributed_pytorch_from_scratch_b200 — a B200-native (sm_100a) tensor-parallel transformer training framework.

Layout
    ops/       ctypes bindings + autograd wrappers around libdpfs_b200.so (hand-written tcgen05/TMA kernels)
    parallel/  process-group manager, communication autograd ops, symmetric-memory runtime
    models/    Column/Row/Vocab-parallel layers, RMSNorm, the LLaMA-style Transformer and its vanilla oracle
    utils/     seeding, distributed bootstrap, checkpoints, data, timers

The reference's flat import paths (``models.layers``, ``process_manager``, ``utils``, ``constants`` ...) are kept
as thin shims at the repository root.
"""

__version__ = "0.1.0"
