#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python tools/bench_attention.py --out gpurun_out/attention_v2.json > gpurun_out/attn_v2.log 2>&1; echo "attn_v2 rc=$?" >> gpurun_out/rc.txt
LLMCTL_FLASH_FWD_V1=1 timeout 300 python tools/bench_attention.py --out gpurun_out/attention_v1.json > gpurun_out/attn_v1.log 2>&1; echo "attn_v1 rc=$?" >> gpurun_out/rc.txt
timeout 200 python bin/llmctl bench kernels --kv-cache --out gpurun_out/bench_paged.json > gpurun_out/bench_paged.log 2>&1; echo "bench_paged rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_serving_gpu.py -m gpu -q --timeout 120 --timeout-method=thread > gpurun_out/t_serving.log 2>&1; echo "serving rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
python - <<'PY'
import json
for f in ("gpurun_out/attention_v2.json","gpurun_out/attention_v1.json"):
    try:
        for r in json.load(open(f)):
            print(f[-7:-5], r["shape"], "native fwd %.3f (%.0f TF) f+b %.3f | cudnn fwd %.3f f+b %.3f | x%.2f x%.2f" % (r["native"]["fwd_ms"], r["native"]["fwd_tflops"], r["native"]["fwd_bwd_ms"], r["cudnn"]["fwd_ms"], r["cudnn"]["fwd_bwd_ms"], r["fwd_speedup_vs_cudnn"], r["fwd_bwd_speedup_vs_cudnn"]))
    except Exception as e: print(f, e)
for r in json.load(open("gpurun_out/bench_paged.json")): print(r["kernel"], r["native_ms"], r.get("achieved_hbm_gbs"))
PY
tail -n 5 gpurun_out/attn_v2.log | cut -c1-300; tail -n 3 gpurun_out/t_serving.log
