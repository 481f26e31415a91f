#!/bin/bash
# Round-2 call I (2 GPUs): first hardware run of the one-kernel GEMM + hand-off; TP=2 bench with it; TP=1 bench (SwiGLU rewrite)
mkdir -p gpurun_out/r2i
O=gpurun_out/r2i
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "gemm_handoff and 2" > $O/test_gh.log 2>&1; rc=$?; echo "test_gh=$rc"; tail -15 $O/test_gh.log | cut -c1-300
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_transformer.py -m gpu -x -q > $O/test_1gpu.log 2>&1; echo "test_1gpu=$?"; tail -3 $O/test_1gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
if [ $rc -eq 0 ]; then
DPFS_SP_FUSED_GEMM=1 timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench2_fusedgemm.log 2>&1; echo "bench2_fusedgemm=$?"; grep '^{' $O/bench2_fusedgemm.log | cut -c1-300
fi
timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench2.log 2>&1; echo "bench2=$?"; grep '^{' $O/bench2.log | cut -c1-300
timeout 240 python bench.py --steps 20 --warmup 5 > $O/bench1.log 2>&1; echo "bench1=$?"; grep '^{' $O/bench1.log | cut -c1-300
timeout 200 python tools/gpu_step_profile.py --tag _r2i > $O/prof1.log 2>&1; cp gpurun_out/step_profile_tp1_r2i.md $O/ 2>/dev/null; sed -n 5,22p gpurun_out/step_profile_tp1_r2i.md
