#!/bin/bash
# Round-2 call F (1 GPU): first hardware run of the opt-in attention kernels + single-GPU test suite
mkdir -p gpurun_out/r2f
O=gpurun_out/r2f
timeout 150 python tools/gpu_attn_debug.py bwd > $O/attn_base.log 2>&1; echo "base=$?"; grep -E "FAIL|time" $O/attn_base.log
DPFS_ATTN_FWD_PP=1 timeout 150 python tools/gpu_attn_debug.py fwd > $O/attn_pp.log 2>&1; echo "pp=$?"; grep -E "FAIL|time|Error" $O/attn_pp.log | head
DPFS_ATTN_BWD_V2=1 timeout 150 python tools/gpu_attn_debug.py bwd > $O/attn_v2.log 2>&1; echo "v2=$?"; grep -E "FAIL|time|Error" $O/attn_v2.log | head
nvidia-smi --query-gpu=name,memory.used --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > $O/tests1.log 2>&1; echo "tests1=$?"
tail -4 $O/tests1.log
