#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "flash" --timeout 60 --timeout-method=thread > gpurun_out/t_flash.log 2>&1; echo "flash rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_engine_gpu.py -m gpu -q --timeout 120 --timeout-method=thread > gpurun_out/t_engine.log 2>&1; echo "engine rc=$?" >> gpurun_out/rc.txt
for c in tp_kernels zero3_native tp_sp_engine; do
  LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py $c > gpurun_out/mgpu_$c.log 2>&1; echo "mgpu_$c rc=$?" >> gpurun_out/rc.txt
done
timeout 600 python bench.py --steps 6 --warmup 3 --no-e2e > gpurun_out/bench_native.log 2>&1; echo "bench_native rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 6 --warmup 3 --no-e2e --attention sdpa > gpurun_out/bench_sdpa.log 2>&1; echo "bench_sdpa rc=$?" >> gpurun_out/rc.txt
timeout 200 python bin/llmctl bench kernels --attention --kv-cache --out gpurun_out/bench_attn.json > gpurun_out/bench_attn.log 2>&1; echo "bench_attn rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 8 gpurun_out/t_flash.log | cut -c1-300; tail -n 4 gpurun_out/t_engine.log | cut -c1-300
for c in tp_kernels zero3_native tp_sp_engine; do echo "== $c"; grep -E "MGPU_RESULT|AssertionError|Error" gpurun_out/mgpu_$c.log | tail -n 3 | cut -c1-1500; done
tail -n 1 gpurun_out/bench_native.log | cut -c1-300; tail -n 1 gpurun_out/bench_sdpa.log | cut -c1-300
grep -E "attention|paged" -A3 gpurun_out/bench_attn.log | head -20
