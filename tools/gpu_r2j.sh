#!/bin/bash
mkdir -p gpurun_out/r2j
python tools/gpu_profile_ew.py > gpurun_out/r2j/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"swiglu|rmsnorm_bwd" -s 4 -c 3 -o gpurun_out/r2j/ew python tools/gpu_profile_ew.py > gpurun_out/r2j/ncu.log 2>&1; echo "ncu=$?"; tail -5 gpurun_out/r2j/ncu.log
