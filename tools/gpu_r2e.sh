#!/bin/bash
# Round-2 call E (8 GPUs): hand-off micro-benchmark + headline bench
mkdir -p gpurun_out/r2e
O=gpurun_out/r2e
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 200 $TR tools/bench_handoff.py > $O/handoff8.log 2>&1; echo "handoff8=$?"
cp gpurun_out/handoff_tp8.md $O/ 2>/dev/null; tail -60 $O/handoff8.log
timeout 200 $TR bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_default.log 2>&1; echo "bench=$?"
grep '^{' $O/bench_default.log | tail -1
