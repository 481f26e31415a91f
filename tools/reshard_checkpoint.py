"""Move a checkpoint between tensor-parallel sizes (the reference can only reload a checkpoint at the TP size that wrote
it: per-rank shard files, reference train.py:121-125 / test.py:49-55).

    python tools/reshard_checkpoint.py --src ckpt_tp2 --dst ckpt_tp8 --src_tp 2 --dst_tp 8 [--step 20000]

Reads `tprank-{r}_iter-{step}_loss-*.pth` for r < src_tp, merges the shards (column-parallel weights/biases and the
embedding along dim 0, row-parallel weights along dim 1, replicated tensors from rank 0) and writes dst_tp shard files
with the same naming.  Without --step the newest iteration present for rank 0 is used.  CPU only.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from distributed_pytorch_from_scratch_b200.utils.checkpoint import (CKPT_RE, list_checkpoints,  # noqa: E402
                                                                   reshard_checkpoint)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--src", required=True, help="directory holding the source shards")
    ap.add_argument("--dst", required=True, help="directory to write the new shards into")
    ap.add_argument("--src_tp", type=int, required=True)
    ap.add_argument("--dst_tp", type=int, required=True)
    ap.add_argument("--step", type=int, default=None, help="iteration to convert (default: newest)")
    args = ap.parse_args(argv)
    step = args.step
    if step is None:
        ck = list_checkpoints(args.src, 0)
        if not ck:
            ap.error(f"no tprank-0_iter-*_loss-*.pth in {args.src}")
        step = int(CKPT_RE.search(os.path.basename(ck[-1])).group(2))
    for p in reshard_checkpoint(args.src, args.dst, step, args.src_tp, args.dst_tp):
        print(p)
    return 0


if __name__ == "__main__":
    sys.exit(main())
