#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "flash" --timeout 60 --timeout-method=thread > gpurun_out/t_flash.log 2>&1; echo "flash rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py watchdog > gpurun_out/mgpu_watchdog.log 2>&1; echo "mgpu_watchdog rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/bench_attention.py --out gpurun_out/attention_v2.json > gpurun_out/attn_v2.log 2>&1; echo "attn_v2 rc=$?" >> gpurun_out/rc.txt
LLMCTL_FLASH_FWD_V1=1 timeout 300 python tools/bench_attention.py --out gpurun_out/attention_v1.json > gpurun_out/attn_v1.log 2>&1; echo "attn_v1 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 12 gpurun_out/t_flash.log | cut -c1-300
grep -E "MGPU_RESULT|Error|assert" gpurun_out/mgpu_watchdog.log | tail -n 4 | cut -c1-800
echo "== v2"; python - <<'PY'
import json
for f in ("gpurun_out/attention_v2.json","gpurun_out/attention_v1.json"):
    try:
        for r in json.load(open(f)):
            print(f[-7:-5], r["shape"], "native fwd %.3f (%.0f TF) f+b %.3f | cudnn fwd %.3f f+b %.3f | x%.2f x%.2f" % (r["native"]["fwd_ms"], r["native"]["fwd_tflops"], r["native"]["fwd_bwd_ms"], r["cudnn"]["fwd_ms"], r["cudnn"]["fwd_bwd_ms"], r["fwd_speedup_vs_cudnn"], r["fwd_bwd_speedup_vs_cudnn"]))
    except Exception as e: print(f, e)
PY
tail -n 5 gpurun_out/attn_v2.log | cut -c1-300
