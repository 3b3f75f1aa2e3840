#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_serving_gpu.py -m gpu -q --timeout 120 --timeout-method=thread > gpurun_out/t_serving.log 2>&1; echo "serving rc=$?" >> gpurun_out/rc.txt
for c in tp_kernels tp_sp_engine; do
  LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py $c > gpurun_out/mgpu_$c.log 2>&1; echo "mgpu_$c rc=$?" >> gpurun_out/rc.txt
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2.log 2>&1; echo "bench_n2 rc=$?" >> gpurun_out/rc.txt
timeout 900 python bench.py --impl reference --steps 6 --warmup 3 > gpurun_out/bench_ref_n1.log 2>&1; echo "bench_ref rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
tail -n 4 gpurun_out/t_serving.log
for c in tp_kernels tp_sp_engine; do echo "== $c"; grep -E "MGPU_RESULT|AssertionError" gpurun_out/mgpu_$c.log | tail -n 3 | cut -c1-1800; done
tail -n 2 gpurun_out/bench_n2.log | cut -c1-1200
tail -n 4 gpurun_out/bench_ref_n1.log | cut -c1-1500
