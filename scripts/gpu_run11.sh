#!/bin/bash
# 8-GPU validation: fused kernels at tp=8, K7 at dp=8, scaling bench N=4,8 (+ reference arm at 8), comms.
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
N=${NGPU:-8}
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
for c in probe tp_kernels zero_native; do
  LLMCTL_DEBUG=1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py $c > gpurun_out/mgpu${N}_$c.log 2>&1; echo "mgpu${N}_$c rc=$?" >> gpurun_out/rc.txt
done
for n in 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$n --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $n --steps 6 --warmup 3 > gpurun_out/bench_n$n.log 2>&1; echo "bench_n$n rc=$?" >> gpurun_out/rc.txt
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29614 bench.py --impl reference --gpus 8 --steps 6 --warmup 3 > gpurun_out/bench_ref_n8.log 2>&1; echo "bench_ref_n8 rc=$?" >> gpurun_out/rc.txt
for p in allreduce allgather reducescatter; do
  for sz in 1MB 1GB; do
    timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node=$N --master-addr 127.0.0.1 --master-port 29613 bin/llmctl bench comms --pattern $p --size $sz --out gpurun_out/comms${N}_${p}_${sz}.json > gpurun_out/comms${N}_${p}_${sz}.log 2>&1; echo "comms $p $sz rc=$?" >> gpurun_out/rc.txt
  done
done
cat gpurun_out/rc.txt
for c in probe tp_kernels zero_native; do echo "== $c"; grep -E "MGPU_RESULT|AssertionError|Error" gpurun_out/mgpu${N}_$c.log | tail -n 3 | cut -c1-1500; done
for n in 4 8; do tail -n 1 gpurun_out/bench_n$n.log | cut -c1-330; done
tail -n 1 gpurun_out/bench_ref_n8.log | cut -c1-400
for f in gpurun_out/comms${N}_*.json; do python -c "import json,sys; d=json.load(open('$f')); print(d['pattern'], d['bytes'], {k:(round(v['time_us'],1), round(v.get('busbw_gbps',0),1)) for k,v in d.items() if isinstance(v,dict)})"; done
