#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
run() { name=$1; shift; timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 2 --steps 4 --warmup 3 --no-e2e --model gpt-1b "$@" > gpurun_out/cfg_$name.log 2>&1; echo "$name rc=$?" >> gpurun_out/rc.txt; tail -n 1 gpurun_out/cfg_$name.log | cut -c1-400; }
run tp2sp --tp 2 --sp
run zero3 --zero-stage 3
run pp2 --pp 2
run tp2sp_lib --tp 2 --sp --impl baseline
cat gpurun_out/rc.txt
grep -h -E "Error|error" gpurun_out/cfg_*.log | sort | uniq -c | head -8 | cut -c1-300
