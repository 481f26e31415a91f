#!/bin/bash
# Round-2 call H (2 GPUs): complete -m gpu suite (single-GPU + TP=2 + full-depth grids), TP=2 bench, TP=1 bench
mkdir -p gpurun_out/r2h
O=gpurun_out/r2h
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > $O/tests_all.log 2>&1; echo "tests=$?"
tail -25 $O/tests_all.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench2.log 2>&1; echo "bench2=$?"; grep '^{' $O/bench2.log | cut -c1-300
DPFS_SP_SIDE_WGRAD=0 timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench2_noside.log 2>&1; echo "bench2_noside=$?"; grep '^{' $O/bench2_noside.log | cut -c1-300
timeout 240 python bench.py --steps 20 --warmup 5 > $O/bench1.log 2>&1; echo "bench1=$?"; grep '^{' $O/bench1.log | cut -c1-300
timeout 200 python tools/gpu_step_profile.py --tag _r2h > $O/prof1.log 2>&1; cp gpurun_out/step_profile_tp1_r2h.md $O/ 2>/dev/null; head -30 gpurun_out/step_profile_tp1_r2h.md
