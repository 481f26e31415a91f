#!/bin/bash
# Round-2 call P (1 GPU): SwiGLU epilogue GEMMs, single-GPU suite, TP=1 bench and profile
mkdir -p gpurun_out/r2p
O=gpurun_out/r2p
timeout 120 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "swiglu" > $O/test_swiglu.log 2>&1; rc=$?; echo "test_swiglu=$rc"; tail -12 $O/test_swiglu.log | cut -c1-220
if [ $rc -ne 0 ]; then exit 1; fi
timeout 240 python bench.py --steps 20 --warmup 5 > $O/bench1.log 2>&1; echo "bench1=$?"; grep '^{' $O/bench1.log | cut -c150-250
timeout 200 python tools/gpu_step_profile.py --tag _r2p > $O/prof1.log 2>&1; cp gpurun_out/step_profile_tp1_r2p.md $O/ 2>/dev/null; sed -n 5,16p gpurun_out/step_profile_tp1_r2p.md
