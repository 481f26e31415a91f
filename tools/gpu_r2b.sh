#!/bin/bash
# Round-2 call B (8 GPUs): TP=8 validation of the fused reduce-scatter / push all-gather kernels, step breakdowns, benches
mkdir -p gpurun_out/r2b
O=gpurun_out/r2b
timeout 400 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "(reduce_scatter and 8) or (push and 8) or (sequence_parallel and fused and 8)" > $O/tests8.log 2>&1; echo "tests8=$?"
tail -4 $O/tests8.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 200 $TR tools/gpu_step_profile.py --sp on --tag _sp_split > $O/prof_split.log 2>&1; echo "prof_split=$?"
DPFS_FUSE_RS=1 DPFS_FUSE_AG=1 timeout 200 $TR tools/gpu_step_profile.py --sp on --tag _sp_rs_ag > $O/prof_rs_ag.log 2>&1; echo "prof_rs_ag=$?"
cp gpurun_out/step_profile_tp8*.md $O/ 2>/dev/null
rb() { tag=$1; shift; env "$@" timeout 200 $TR bench.py --gpus 8 --steps 10 --warmup 3 $EXTRA > $O/bench_$tag.log 2>&1; echo "$tag=$?"; grep '^{' $O/bench_$tag.log | python -c "import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d.get('ms_per_step'), d.get('final_loss'), d['config'].get('parallelism'), d.get('kernels_per_step'))"; }
EXTRA="" rb default A=1
EXTRA="" rb rs_ag DPFS_FUSE_RS=1 DPFS_FUSE_AG=1
EXTRA="" rb ag DPFS_FUSE_AG=1
EXTRA="--sp off" rb nosp_nvls A=1
EXTRA="--sp off" rb nosp_p2p_fusedar DPFS_AR_MODE=p2p
