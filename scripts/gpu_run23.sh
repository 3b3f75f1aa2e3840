#!/bin/bash
# 8-GPU validation: fused TP kernels (K1 copy pipeline), TP decode with K11, ZeRO-3 peer gather, training bench
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
for c in tp_kernels tp_serving zero3_native; do
  LLMCTL_DEBUG=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py $c > gpurun_out/mgpu8_$c.log 2>&1; echo "mgpu8_$c rc=$?" >> gpurun_out/rc.txt
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 6 --warmup 3 > gpurun_out/bench_n8.log 2>&1; echo "bench8 rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt
for c in tp_kernels tp_serving zero3_native; do echo "== $c"; grep -E "MGPU_RESULT|AssertionError|Error" gpurun_out/mgpu8_$c.log | tail -n 2 | cut -c1-1500; done
tail -n 1 gpurun_out/bench_n8.log | cut -c1-1200
