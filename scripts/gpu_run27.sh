#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node=8 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_worker.py tp_kernels > gpurun_out/mgpu8_tp_kernels.log 2>&1; echo "mgpu8_tp_kernels rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; grep -E "MGPU_RESULT|AssertionError|Error" gpurun_out/mgpu8_tp_kernels.log | tail -n 2 | cut -c1-1800
