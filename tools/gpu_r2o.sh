#!/bin/bash
# Round-2 call O (1 GPU): compute-sanitizer memcheck over the kernel unit tests of the memory-bound kernels + small GEMMs
mkdir -p gpurun_out/r2o
O=gpurun_out/r2o
SEL="rmsnorm or layernorm or swiglu_gelu or rope or colsum or cross_entropy or fused_adam_kernel or (gemm_tcgen05 and 128-128-64) or (flash_attention and 1-128-1-64)"
timeout 120 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "$SEL" > $O/plain.log 2>&1; echo "plain=$?"; tail -2 $O/plain.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "$SEL" > $O/memcheck.log 2>&1; echo "memcheck=$?"; grep -E "ERROR SUMMARY|passed|failed|Invalid|error" $O/memcheck.log | head -12
