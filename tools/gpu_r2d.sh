#!/bin/bash
# Round-2 call D (2 GPUs): all TP=2 tests, hand-off micro-benchmark, TP=2 benches (non-SP fused all-reduce vs SP hand-offs)
mkdir -p gpurun_out/r2d
O=gpurun_out/r2d
timeout 600 python -m pytest tests/test_gpu_tp.py -m gpu -x -q > $O/tests2.log 2>&1; echo "tests2=$?"
tail -5 $O/tests2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
timeout 300 $TR tools/bench_handoff.py > $O/handoff2.log 2>&1; echo "handoff2=$?"
cp gpurun_out/handoff_tp2.md $O/ 2>/dev/null; tail -50 $O/handoff2.log
rb() { tag=$1; shift; env "$@" timeout 240 $TR bench.py --gpus 2 --steps 10 --warmup 3 $EXTRA > $O/bench_$tag.log 2>&1; echo "$tag=$?"; grep '^{' $O/bench_$tag.log | python -c "import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d.get('ms_per_step'), d.get('final_loss'), d['config'].get('parallelism'), d.get('kernels_per_step'), d.get('parity'))"; }
EXTRA="" rb default A=1
EXTRA="--sp on" rb sp_handoff A=1
