#!/bin/bash
# Round-2 call W (2 GPUs): the real training entrypoint on the hand-off path (ragged batches: several graph shapes, then eager)
mkdir -p gpurun_out/r2w
timeout 240 python train.py --tp_size 2 --bf16 --cuda_graph --sequence_parallel --synthetic_data --save_dir /tmp/ck_r2w --max_steps 60 --warmup_steps 5 --log_interval 20 --save_interval 1000 --profile --master_port 29566 > gpurun_out/r2w/train_tp2_sp.log 2>&1; echo "train=$?"; grep -E "profile|Avg Loss|finished|Error|error" gpurun_out/r2w/train_tp2_sp.log | head -12 | cut -c1-220
