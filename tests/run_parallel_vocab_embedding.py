"""Script-style entry point with the reference's CLI (its tests/test_parallel_vocab_embedding.py:1-178 is run as
`python tests/test_parallel_vocab_embedding.py --tp_size N`): spawns N ranks (`--backend gloo`: the CPU-sized grids of the pytest suite; `--backend nccl`: the reference's full grid and 1000-step histories on GPUs, fp32, like the reference's own scripts) and runs the same
assertions as the pytest suite.  Unlike the reference, a failing assertion makes the process exit non-zero."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from dist_utils import run_distributed  # noqa: E402
from test_parallel_layers import _emb_worker  # noqa: E402

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--master_addr", type=str, default="127.0.0.1")
    ap.add_argument("--master_port", type=str, default="25555")
    ap.add_argument("--random_seed", type=int, default=42)
    ap.add_argument("--tp_size", type=int, default=2)
    ap.add_argument("--backend", type=str, default="gloo", choices=["gloo", "nccl"])
    a = ap.parse_args()
    if a.backend == "nccl":      # one process per GPU, NCCL, the reference's grid at full depth
        from test_gpu_layer_grids import _embedding_worker  # noqa: E402
        run_distributed(_embedding_worker, a.tp_size, int(a.master_port), backend="nccl", args=())
    else:
        run_distributed(_emb_worker, a.tp_size, int(a.master_port), backend=a.backend)
    print(f"OK: parallel_vocab_embedding at TP={a.tp_size} ({a.backend})")
