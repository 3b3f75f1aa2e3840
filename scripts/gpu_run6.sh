#!/bin/bash
# 2-GPU run: serving tests (after the shuffle fix), symmetric memory probe, fused kernels, K7.
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_serving_gpu.py -m gpu -q -x --timeout 120 --timeout-method=thread > gpurun_out/t_serving.log 2>&1; echo "serving rc=$?" >> gpurun_out/rc.txt
for c in probe tp_kernels zero_native tp_sp_engine; do
  LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py $c > gpurun_out/mgpu_$c.log 2>&1; echo "mgpu_$c rc=$?" >> gpurun_out/rc.txt
done
timeout 300 python bin/llmctl bench e2e --model gpt-1b --prompt-length 512 --gen-length 64 --num-requests 8 --out gpurun_out/serve_1b.json > gpurun_out/serve_1b.log 2>&1; echo "serve1b rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -n 6 gpurun_out/t_serving.log
for c in probe tp_kernels zero_native tp_sp_engine; do echo "== $c"; grep -E "MGPU_RESULT|Error|error|assert" gpurun_out/mgpu_$c.log | tail -n 6 | cut -c1-1500; done
tail -n 12 gpurun_out/serve_1b.log
