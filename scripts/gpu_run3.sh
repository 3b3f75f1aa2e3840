#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -q -x --timeout 300 --timeout-method=thread > gpurun_out/t_engine.log 2>&1; echo "engine rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py --layers 4 --out gpurun_out/profile_native.txt > gpurun_out/profile_native.log 2>&1; echo "prof rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py --layers 4 --gemm cublas --out gpurun_out/profile_cublas.txt > gpurun_out/profile_cublas.log 2>&1; echo "prof2 rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --no-e2e > gpurun_out/bench_native.log 2>&1; echo "bench_native rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -n 12 gpurun_out/smoke.log; tail -n 15 gpurun_out/t_engine.log
cat gpurun_out/profile_native.txt | head -45
tail -n 1 gpurun_out/bench_native.log | cut -c1-300
