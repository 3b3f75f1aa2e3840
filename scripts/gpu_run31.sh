#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 100 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "skinny" --timeout 30 --timeout-method=thread > gpurun_out/t_skinny.log 2>&1; echo "skinny rc=$?" >> gpurun_out/rc.txt
timeout 200 python tools/bench_skinny.py --out gpurun_out/skinny_gemm.json > gpurun_out/bench_skinny.log 2>&1; echo "bench_skinny rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_decode.py gpt-7b graph > gpurun_out/pd_graph.log 2>&1; echo "graph rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 3 gpurun_out/t_skinny.log | cut -c1-200
python - <<'PY'
import json
for r in json.load(open("gpurun_out/skinny_gemm.json")):
    print(r["layer"], r["weight_MB"], "MB | skinny %.1f us %.2f TB/s | general %.1f us %.2f | cublas %.1f us %.2f" % (1e3*r["skinny_ms"], r["skinny_TBps"], 1e3*r["general_ms"], r["general_TBps"], 1e3*r["cublas_ms"], r["cublas_TBps"]))
PY
grep round gpurun_out/pd_graph.log | cut -c1-200
