#!/bin/bash
# Round-2 call M (4 GPUs): TP=4 benches (one-kernel GEMM + hand-off vs split) and the TP=4 tests of the new kernels
mkdir -p gpurun_out/r2m
O=gpurun_out/r2m
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544"
timeout 200 $TR bench.py --gpus 4 --steps 10 --warmup 3 > $O/bench4.log 2>&1; echo "bench4=$?"; grep '^{' $O/bench4.log | cut -c150-250
DPFS_SP_FUSED_GEMM=0 timeout 200 $TR bench.py --gpus 4 --steps 10 --warmup 3 > $O/bench4_split.log 2>&1; echo "bench4_split=$?"; grep '^{' $O/bench4_split.log | cut -c150-250
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -q -k "(handoff and 4) or (layer_classes and 4)" > $O/tests4.log 2>&1; echo "tests4=$?"; tail -3 $O/tests4.log | cut -c1-200
