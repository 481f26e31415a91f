#!/bin/bash
# Round-2 call U (2 GPUs): quick check of the SP paths after the zone-sizing change
mkdir -p gpurun_out/r2u
O=gpurun_out/r2u
timeout 200 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "(sequence_parallel and fused and 2) or (one_kernel_gemm and 2)" > $O/tests2.log 2>&1; echo "tests2=$?"; tail -2 $O/tests2.log | cut -c1-200
