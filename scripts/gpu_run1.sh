#!/bin/bash
# First GPU batch: kernel numerics, smoke, GEMM perf, bench (cuBLAS-GEMM fallback first so a bench line exists).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "not gemm and not linear" --timeout 120 --timeout-method=thread > gpurun_out/t_elem.log 2>&1; echo "elem rc=$?" >> gpurun_out/rc.txt
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "gemm or linear" --timeout 120 --timeout-method=thread > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?" >> gpurun_out/rc.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/bench_gemm.py --out gpurun_out/gemm_bench.json > gpurun_out/gemm_bench.log 2>&1; echo "gemmbench rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 4 --warmup 3 --gemm cublas > gpurun_out/bench_cublasgemm.log 2>&1; echo "bench_cublas rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_native.log 2>&1; echo "bench_native rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 4 --warmup 3 --impl baseline --no-e2e > gpurun_out/bench_baseline.log 2>&1; echo "bench_baseline rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -5 gpurun_out/t_elem.log gpurun_out/t_gemm.log gpurun_out/smoke.log
tail -3 gpurun_out/bench_*.log
