#!/bin/bash
# BASELINE.md §4 configurations at 8 GPUs (GPT-7B TP=8+SP, GPT-7B ZeRO-3 DP=8, GPT-13B TP=4 x PP=2)
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
run() { name=$1; shift; timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 8 --steps 4 --warmup 3 --no-e2e "$@" > gpurun_out/cfg8_$name.log 2>&1; echo "$name rc=$?" >> gpurun_out/rc.txt; tail -n 1 gpurun_out/cfg8_$name.log | cut -c1-900; }
run tp8sp --tp 8 --sp
run zero3 --zero-stage 3
run 13b_tp4pp2 --model gpt-13b --tp 4 --pp 2 --micro-batch 4 --seq-len 2048
cat gpurun_out/rc.txt
grep -h -E "Error" gpurun_out/cfg8_*.log | sort | uniq -c | head -6 | cut -c1-300
