#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "flash" --timeout 60 --timeout-method=thread > gpurun_out/t_flash.log 2>&1; echo "flash rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py tp_kernels > gpurun_out/mgpu_tp_kernels.log 2>&1; echo "mgpu_tp_kernels rc=$?" >> gpurun_out/rc.txt
for p in allreduce allgather reducescatter; do
  for sz in 1MB 64MB 1GB; do
    timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29613 bin/llmctl bench comms --pattern $p --size $sz --out gpurun_out/comms_${p}_${sz}.json > gpurun_out/comms_${p}_${sz}.log 2>&1; echo "comms $p $sz rc=$?" >> gpurun_out/rc.txt
  done
done
timeout 200 python bin/llmctl bench kernels --attention --out gpurun_out/bench_attn.json > gpurun_out/bench_attn.log 2>&1; echo "bench_attn rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 6 gpurun_out/t_flash.log | cut -c1-300
grep -E "MGPU_RESULT|AssertionError" gpurun_out/mgpu_tp_kernels.log | tail -n 2 | cut -c1-1500
for f in gpurun_out/comms_*.json; do python -c "import json,sys; d=json.load(open('$f')); print(d['pattern'], d['bytes'], {k:(round(v['time_us'],1), round(v.get('busbw_gbps',0),1)) for k,v in d.items() if isinstance(v,dict)})"; done
tail -n 12 gpurun_out/bench_attn.log
