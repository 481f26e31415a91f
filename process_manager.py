"""Reference-compatible import path (``import process_manager as pm``; reference process_manager.py:1-25).

This module *is* distributed_pytorch_from_scratch_b200.parallel.process_manager, so ``pm.pgm`` always reflects the
live manager set by ``pm.init_pgm``.
"""
import sys

from distributed_pytorch_from_scratch_b200.parallel import process_manager as _impl

sys.modules[__name__] = _impl
