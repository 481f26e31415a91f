#!/bin/bash
# Round-2 call C (1 GPU): full single-GPU test suite + headline bench
mkdir -p gpurun_out/r2c
O=gpurun_out/r2c
timeout 900 python -m pytest tests -m gpu -x -q > $O/tests1.log 2>&1; echo "tests1=$?"
tail -6 $O/tests1.log
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench1.log 2>&1; echo "bench1=$?"
grep '^{' $O/bench1.log | tail -1
