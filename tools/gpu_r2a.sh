#!/bin/bash
# Round-2 call A (2 GPUs): first hardware run of the opt-in fused reduce-scatter / push all-gather kernels + TP=2 benches
mkdir -p gpurun_out/r2a
O=gpurun_out/r2a
export DPFS_TEST_EXPERIMENTAL=1
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "reduce_scatter" > $O/rs.log 2>&1; echo "rs=$?"
tail -5 $O/rs.log
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -x -q -k "push" > $O/push.log 2>&1; echo "push=$?"
tail -5 $O/push.log
rb() { tag=$1; shift; env "$@" timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 10 --warmup 3 $EXTRA > $O/bench_$tag.log 2>&1; echo "$tag=$?"; grep '^{' $O/bench_$tag.log | python -c "import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d.get('ms_per_step'), d.get('final_loss'), d['config'].get('parallelism'), d.get('kernels_per_step'))"; }
EXTRA="" rb default A=1
EXTRA="--sp on" rb sp_p2p A=1
EXTRA="--sp on" rb sp_nvls DPFS_AR_MODE=nvls
EXTRA="--sp on" rb sp_nvls_rs DPFS_AR_MODE=nvls DPFS_FUSE_RS=1
EXTRA="--sp on" rb sp_nvls_rs_ag DPFS_AR_MODE=nvls DPFS_FUSE_RS=1 DPFS_FUSE_AG=1
EXTRA="--sp on" rb sp_p2p_rs DPFS_FUSE_RS=1
