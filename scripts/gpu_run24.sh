#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 100 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "flash" --timeout 30 --timeout-method=thread > gpurun_out/t_flash.log 2>&1; echo "flash rc=$?" >> gpurun_out/rc.txt
timeout 100 python tools/bench_attention.py --out gpurun_out/attention_v4.json > gpurun_out/attn_v4.log 2>&1; echo "attn rc=$?" >> gpurun_out/rc.txt
echo skip
cat gpurun_out/rc.txt; tail -n 4 gpurun_out/t_flash.log | cut -c1-300
python - <<'PY'
import json
for r in json.load(open("gpurun_out/attention_v4.json")):
    print(r["shape"], "native fwd %.3f (%.0f TF) f+b %.3f | cudnn fwd %.3f f+b %.3f | x%.2f x%.2f" % (r["native"]["fwd_ms"], r["native"]["fwd_tflops"], r["native"]["fwd_bwd_ms"], r["cudnn"]["fwd_ms"], r["cudnn"]["fwd_bwd_ms"], r["fwd_speedup_vs_cudnn"], r["fwd_bwd_speedup_vs_cudnn"]))
PY
grep -E "MGPU_RESULT|Error|assert" gpurun_out/mgpu_tp_kernels.log | tail -n 3 | cut -c1-1500
