#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 900 python -m pytest tests -m gpu -q --timeout 240 --timeout-method=thread > gpurun_out/t_all_gpu.log 2>&1; echo "pytest_gpu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 15 gpurun_out/t_all_gpu.log | cut -c1-300
