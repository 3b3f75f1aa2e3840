#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "flash" --timeout 60 --timeout-method=thread > gpurun_out/t_flash.log 2>&1; echo "flash rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/bench_attention.py --out gpurun_out/attention_v3.json > gpurun_out/attn_v3.log 2>&1; echo "attn rc=$?" >> gpurun_out/rc.txt
timeout 600 python bin/llmctl bench e2e --gen-length 128 --out gpurun_out/bench_e2e_graph.json > gpurun_out/bench_e2e_graph.log 2>&1; echo "e2e rc=$?" >> gpurun_out/rc.txt
timeout 600 python bin/llmctl bench e2e --gen-length 128 --no-cuda-graphs --out gpurun_out/bench_e2e_eager.json > gpurun_out/bench_e2e_eager.log 2>&1; echo "e2e eager rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 6 gpurun_out/t_flash.log | cut -c1-300
python - <<'PY'
import json
for r in json.load(open("gpurun_out/attention_v3.json")):
    print(r["shape"], "native fwd %.3f (%.0f TF) f+b %.3f | cudnn fwd %.3f f+b %.3f | x%.2f x%.2f" % (r["native"]["fwd_ms"], r["native"]["fwd_tflops"], r["native"]["fwd_bwd_ms"], r["cudnn"]["fwd_ms"], r["cudnn"]["fwd_bwd_ms"], r["fwd_speedup_vs_cudnn"], r["fwd_bwd_speedup_vs_cudnn"]))
for f in ("graph","eager"):
    print(f, json.load(open(f"gpurun_out/bench_e2e_{f}.json")))
PY
