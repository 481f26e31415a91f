#!/bin/bash
# Round-2 call G (8 GPUs): hand-off kernels at TP=8 (stand-alone and fused with the GEMM), micro-benchmark, per-rank
# step profile, benches with / without the one-kernel GEMM + hand-off
mkdir -p gpurun_out/r2g
O=gpurun_out/r2g
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "(handoff and 8) or (layer_classes and 8)" > $O/tests8.log 2>&1; rc=$?; echo "tests8=$rc"; tail -4 $O/tests8.log | cut -c1-250
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 200 $TR tools/bench_handoff.py > $O/handoff8.log 2>&1; echo "handoff8=$?"
cp gpurun_out/handoff_tp8.md $O/ 2>/dev/null; grep -E "^\|" $O/handoff8.log | grep -E "grid-wide, 1 fence|both|none \| auto" | head -30
DPFS_SP_FUSED_GEMM=0 timeout 200 $TR bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_split.log 2>&1; echo "bench_split=$?"
grep '^{' $O/bench_split.log | tail -1 | cut -c150-330
if [ $rc -eq 0 ]; then
timeout 200 $TR bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_fused.log 2>&1; echo "bench_fused=$?"
grep '^{' $O/bench_fused.log | tail -1 | cut -c150-330
timeout 200 $TR tools/gpu_step_profile.py --sp on --tag _r2g > $O/prof.log 2>&1; echo "prof=$?"
cp gpurun_out/step_profile_tp8_r2g.md $O/ 2>/dev/null; sed -n 5,16p gpurun_out/step_profile_tp8_r2g.md
fi
