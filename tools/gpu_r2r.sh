#!/bin/bash
# Round-2 call R (2 GPUs): SP model tests + TP=2 bench with the SwiGLU epilogue GEMMs on the hand-off path
mkdir -p gpurun_out/r2r
O=gpurun_out/r2r
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "(sequence_parallel and fused and 2) or (one_kernel and 2) or (fused_transformer and fused and 2)" > $O/tests2.log 2>&1; echo "tests2=$?"; tail -3 $O/tests2.log | cut -c1-200
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
timeout 240 $TR bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench2.log 2>&1; echo "bench2=$?"; grep '^{' $O/bench2.log | cut -c150-250
