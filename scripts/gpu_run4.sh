#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "gemm or linear" --timeout 120 --timeout-method=thread > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 600 python -m pytest tests/test_engine_gpu.py -m gpu -q --timeout 300 --timeout-method=thread > gpurun_out/t_engine.log 2>&1; echo "engine rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/bench_gemm.py --out gpurun_out/gemm_bench.json > gpurun_out/gemm_bench.log 2>&1; echo "gemmbench rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_step.py --layers 4 --out gpurun_out/profile_native.txt > gpurun_out/profile_native.log 2>&1; echo "prof rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --no-e2e > gpurun_out/bench_native.log 2>&1; echo "bench_native rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -n 8 gpurun_out/t_gemm.log; tail -n 4 gpurun_out/smoke.log; tail -n 12 gpurun_out/t_engine.log
python - <<'PY'
import json
for r in json.load(open("gpurun_out/gemm_bench.json")):
    print(r["name"].ljust(16), "cublas %.0f" % r["cublas_tflops"], " ".join(f"{k[7:-7]}={v:.0f}" for k, v in r.items() if k.endswith("_tflops") and k.startswith("native")))
PY
head -n 22 gpurun_out/profile_native.txt | cut -c1-150
tail -n 1 gpurun_out/bench_native.log | cut -c1-200
