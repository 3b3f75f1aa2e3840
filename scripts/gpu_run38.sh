#!/bin/bash
mkdir -p gpurun_out
LLMCTL_DEBUG=1 timeout 400 python -m pytest tests/test_multigpu.py -m gpu -q -x --timeout 200 --timeout-method=thread > gpurun_out/t_mgpu.log 2>&1; echo "rc=$?"; tail -n 4 gpurun_out/t_mgpu.log | cut -c1-200
