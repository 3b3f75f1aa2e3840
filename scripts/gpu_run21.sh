#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python tools/bench_skinny.py --out gpurun_out/skinny_gemm.json > gpurun_out/bench_skinny.log 2>&1; echo "bench_skinny rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py tp_serving > gpurun_out/mgpu_tp_serving.log 2>&1; echo "tp_serving rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
python - <<'PY'
import json
for r in json.load(open("gpurun_out/skinny_gemm.json")):
    print(r["layer"], r["weight_MB"], "MB | skinny %.1f us %.2f TB/s | general %.1f us %.2f | cublas %.1f us %.2f" % (1e3*r["skinny_ms"], r["skinny_TBps"], 1e3*r["general_ms"], r["general_TBps"], 1e3*r["cublas_ms"], r["cublas_TBps"]))
PY
tail -3 gpurun_out/bench_skinny.log | cut -c1-300
grep -E "MGPU_RESULT|Error|assert" gpurun_out/mgpu_tp_serving.log | tail -n 3 | cut -c1-700
