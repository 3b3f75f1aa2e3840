#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "skinny" --timeout 60 --timeout-method=thread > gpurun_out/t_skinny.log 2>&1; echo "skinny rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_serving_gpu.py -m gpu -q --timeout 120 --timeout-method=thread > gpurun_out/t_serving.log 2>&1; echo "serving rc=$?" >> gpurun_out/rc.txt
timeout 200 python tools/bench_skinny.py --out gpurun_out/skinny_gemm.json > gpurun_out/bench_skinny.log 2>&1; echo "bench_skinny rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_decode.py gpt-7b graph > gpurun_out/pd_graph.log 2>&1; echo "graph rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py tp_serving > gpurun_out/mgpu_tp_serving.log 2>&1; echo "tp_serving rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 8 gpurun_out/t_skinny.log | cut -c1-300; tail -n 4 gpurun_out/t_serving.log | cut -c1-300
cat gpurun_out/bench_skinny.log | cut -c1-400; grep round gpurun_out/pd_graph.log | cut -c1-300
grep -E "MGPU_RESULT|Error|assert" gpurun_out/mgpu_tp_serving.log | tail -n 4 | cut -c1-900
