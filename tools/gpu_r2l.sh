#!/bin/bash
# Round-2 call L (2 GPUs): complete -m gpu suite on the current tree, TP=2 benches, RowParallelLinear sweep at TP=2
mkdir -p gpurun_out/r2l
O=gpurun_out/r2l
timeout 1500 python -m pytest tests -m gpu -x -q > $O/tests_all.log 2>&1; echo "tests=$?"
tail -6 $O/tests_all.log | cut -c1-250
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench2.log 2>&1; echo "bench2=$?"; grep '^{' $O/bench2.log | cut -c150-250
DPFS_SP_FUSED_GEMM=0 timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 > $O/bench2_split.log 2>&1; echo "bench2_split=$?"; grep '^{' $O/bench2_split.log | cut -c150-250
timeout 240 python bench.py --steps 20 --warmup 5 > $O/bench1.log 2>&1; echo "bench1=$?"; grep '^{' $O/bench1.log | cut -c150-250
DPFS_GEMM_SWIGLU=0 timeout 240 python bench.py --steps 20 --warmup 5 > $O/bench1_noswiglu.log 2>&1; echo "bench1_noswiglu=$?"; grep '^{' $O/bench1_noswiglu.log | cut -c150-250
timeout 300 $TR tools/bench_rowpar.py --out $O/rowpar_tp2.md > $O/rowpar2.log 2>&1; echo "rowpar2=$?"; tail -3 $O/rowpar2.log | cut -c1-250
