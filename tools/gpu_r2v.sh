#!/bin/bash
mkdir -p gpurun_out/r2v
timeout 200 python tools/gpu_gemm_vs_cublas.py > gpurun_out/r2v/gemm_vs_cublas.log 2>&1; echo "rc=$?"; cp gpurun_out/gemm_vs_cublas.md gpurun_out/r2v/ 2>/dev/null; tail -18 gpurun_out/r2v/gemm_vs_cublas.log | cut -c1-200
