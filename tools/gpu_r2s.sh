#!/bin/bash
# Round-2 call S (2 GPUs): BASELINE.json config 3 (GPT-2-medium shape) at TP=1 and TP=2, both arms
mkdir -p gpurun_out/r2s
O=gpurun_out/r2s
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544"
timeout 200 python bench.py --config gpt2m --steps 10 --warmup 3 > $O/gpt2m_tp1.log 2>&1; echo "ours1=$?"; grep '^{' $O/gpt2m_tp1.log | cut -c150-260
timeout 200 $TR bench.py --config gpt2m --gpus 2 --steps 10 --warmup 3 > $O/gpt2m_tp2.log 2>&1; echo "ours2=$?"; grep '^{' $O/gpt2m_tp2.log | cut -c150-260
timeout 200 $TR bench.py --impl reference --config gpt2m --gpus 2 --steps 5 --warmup 3 > $O/gpt2m_tp2_ref.log 2>&1; echo "ref2=$?"; grep '^{' $O/gpt2m_tp2_ref.log | cut -c1-400
