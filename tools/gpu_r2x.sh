#!/bin/bash
# Round-2 call X (1 GPU): final check of the shipped library -- smoke, -m gpu suite, default bench
mkdir -p gpurun_out/r2x
O=gpurun_out/r2x
timeout 100 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "smoke=$?"; tail -1 $O/smoke.log | cut -c1-120
timeout 300 python -m pytest tests -m gpu -x -q > $O/tests1.log 2>&1; echo "tests1=$?"; tail -1 $O/tests1.log | cut -c1-120
timeout 100 python bench.py --steps 10 --warmup 3 > $O/bench1.log 2>&1; echo "bench1=$?"; grep '^{' $O/bench1.log | cut -c150-240
