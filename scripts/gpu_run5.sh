#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 900 python -m pytest tests -m gpu -q --timeout 300 --timeout-method=thread > gpurun_out/t_all.log 2>&1; echo "tests rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_native.log 2>&1; echo "bench_native rc=$?" >> gpurun_out/rc.txt
timeout 600 python bin/llmctl bench e2e --model gpt-1b --prompt-length 512 --gen-length 64 --num-requests 8 --out gpurun_out/serve_1b.json > gpurun_out/serve_1b.log 2>&1; echo "serve1b rc=$?" >> gpurun_out/rc.txt
timeout 300 python bin/llmctl bench kernels --matmul --norm --rope --optimizer --attention --kv-cache --out gpurun_out/bench_kernels.json > gpurun_out/bench_kernels.log 2>&1; echo "benchk rc=$?" >> gpurun_out/rc.txt
timeout 200 python bin/llmctl hw probe --emit gpurun_out/hw_b200.toml > gpurun_out/hw_probe.log 2>&1; echo "hw rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -n 15 gpurun_out/t_all.log
tail -n 1 gpurun_out/bench_native.log | cut -c1-250
tail -n 12 gpurun_out/serve_1b.log; tail -n 20 gpurun_out/bench_kernels.log
