#!/bin/bash
# Round-2 call K (8 GPUs): TP=8 tests of the hand-off kernels, benches with / without the one-kernel GEMM + hand-off, profile
export DPFS_GEMM_SWIGLU=0   # the SwiGLU-epilogue GEMM has not run on hardware yet: keep it out of this call
mkdir -p gpurun_out/r2k
O=gpurun_out/r2k
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -q -k "(handoff and 8) or (layer_classes and 8)" > $O/tests8.log 2>&1; rc=$?; echo "tests8=$rc"; tail -4 $O/tests8.log | cut -c1-250
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
DPFS_SP_FUSED_GEMM=0 timeout 200 $TR bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_split.log 2>&1; echo "bench_split=$?"
grep '^{' $O/bench_split.log | tail -1 | cut -c150-250
timeout 200 $TR bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_fused.log 2>&1; echo "bench_fused=$?"
grep '^{' $O/bench_fused.log | tail -1 | cut -c150-250
timeout 200 $TR tools/gpu_step_profile.py --sp on --tag _r2k > $O/prof.log 2>&1; echo "prof=$?"
cp gpurun_out/step_profile_tp8_r2k.md $O/ 2>/dev/null; sed -n 5,16p gpurun_out/step_profile_tp8_r2k.md
