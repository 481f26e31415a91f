#!/bin/bash
# Round-2 call N (8 GPUs): benches of record (default = split at TP=8, and the comm-CTA-limited one-kernel form),
# per-rank step profile of the default, RowParallelLinear sweep at TP=8
mkdir -p gpurun_out/r2n
O=gpurun_out/r2n
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 200 $TR bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench8.log 2>&1; echo "bench8=$?"; grep '^{' $O/bench8.log | cut -c150-250
DPFS_SP_FUSED_GEMM=1 timeout 200 $TR bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench8_fused.log 2>&1; echo "bench8_fused=$?"; grep '^{' $O/bench8_fused.log | cut -c150-250
timeout 200 $TR tools/gpu_step_profile.py --sp on --tag _r2n > $O/prof.log 2>&1; echo "prof=$?"
cp gpurun_out/step_profile_tp8_r2n.md $O/ 2>/dev/null; sed -n 5,14p gpurun_out/step_profile_tp8_r2n.md
timeout 200 $TR tools/bench_rowpar.py --out $O/rowpar_tp8.md > $O/rowpar8.log 2>&1; echo "rowpar8=$?"; tail -2 $O/rowpar8.log | cut -c1-200
