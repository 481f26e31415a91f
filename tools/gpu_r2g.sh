#!/bin/bash
# Round-2 call G (8 GPUs): hand-off kernel test with the single-fence grid barrier, micro-benchmark, per-rank step profile, bench
mkdir -p gpurun_out/r2g
O=gpurun_out/r2g
timeout 200 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "handoff and 8" > $O/test_handoff8.log 2>&1; rc=$?; echo "test_handoff8=$rc"; tail -3 $O/test_handoff8.log
if [ $rc -ne 0 ]; then exit 1; fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 200 $TR tools/bench_handoff.py > $O/handoff8.log 2>&1; echo "handoff8=$?"
cp gpurun_out/handoff_tp8.md $O/ 2>/dev/null; grep -E "^\|" $O/handoff8.log | tail -75
timeout 200 $TR tools/gpu_step_profile.py --sp on --tag _r2g > $O/prof.log 2>&1; echo "prof=$?"
cp gpurun_out/step_profile_tp8_r2g.md $O/ 2>/dev/null
DPFS_SP_GRID_CAP=37 timeout 200 $TR bench.py --gpus 8 --steps 10 --warmup 3 > $O/bench_cap37.log 2>&1; echo "bench=$?"
grep '^{' $O/bench_cap37.log | tail -1 | cut -c1-400
