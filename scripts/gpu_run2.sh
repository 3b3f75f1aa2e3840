#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 900 python -m pytest tests/test_kernels_gpu.py -m gpu -q --timeout 120 --timeout-method=thread > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?" >> gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 > gpurun_out/bench_native.log 2>&1; echo "bench_native rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --gemm cublas --no-e2e > gpurun_out/bench_cublasgemm.log 2>&1; echo "bench_cublas rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -n 5 gpurun_out/t_kernels.log gpurun_out/smoke.log
tail -n 2 gpurun_out/bench_native.log gpurun_out/bench_cublasgemm.log
