#!/bin/bash
mkdir -p gpurun_out/r2q
timeout 200 python tools/gpu_step_profile.py --tag _r2q > gpurun_out/r2q/prof1.log 2>&1; cp gpurun_out/step_profile_tp1_r2q.md gpurun_out/r2q/ 2>/dev/null; sed -n 3,24p gpurun_out/step_profile_tp1_r2q.md
