#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
run() { name=$1; shift; timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 2 --steps 4 --warmup 3 --model gpt-1b "$@" > gpurun_out/cfg_$name.log 2>&1; echo "$name rc=$?" >> gpurun_out/rc.txt; tail -n 1 gpurun_out/cfg_$name.log | cut -c1-300; }
run pp2 --pp 2
cat gpurun_out/rc.txt
grep -h -E "Error" gpurun_out/cfg_pp2.log | sort | uniq -c | head -5 | cut -c1-300
