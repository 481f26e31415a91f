#!/bin/bash
# Round-2 call T (1 GPU): what the driver runs at round end on one GPU -- smoke, -m gpu suite, both bench arms
mkdir -p gpurun_out/r2t
O=gpurun_out/r2t
timeout 200 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "smoke=$?"; tail -1 $O/smoke.log | cut -c1-200
timeout 600 python -m pytest tests -m gpu -x -q > $O/tests1.log 2>&1; echo "tests1=$?"; tail -2 $O/tests1.log | cut -c1-200
timeout 240 python bench.py --impl reference --steps 5 --warmup 3 > $O/bench1_ref.log 2>&1; echo "ref1=$?"; grep '^{' $O/bench1_ref.log | cut -c1-330
timeout 240 python bench.py > $O/bench1.log 2>&1; echo "bench1=$?"; grep '^{' $O/bench1.log | cut -c150-250
