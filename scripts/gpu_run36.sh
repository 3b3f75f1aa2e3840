#!/bin/bash
mkdir -p gpurun_out
LLMCTL_DEBUG=1 timeout 200 python -m pytest tests/test_engine_gpu.py tests/test_serving_gpu.py -m gpu -q -x --timeout 120 --timeout-method=thread > gpurun_out/t_sanity.log 2>&1; echo "rc=$?"; tail -n 3 gpurun_out/t_sanity.log | cut -c1-200
