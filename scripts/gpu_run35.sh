#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 150 python -m pytest tests/test_kernels_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x -k "norm or no_alloc" --timeout 60 --timeout-method=thread > gpurun_out/t_norm.log 2>&1; echo "norm rc=$?" >> gpurun_out/rc.txt
timeout 200 python tools/profile_step.py --layers 8 --out gpurun_out/profile_step_8L.txt > gpurun_out/profile_step_8L.log 2>&1; echo "prof rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 3 gpurun_out/t_norm.log | cut -c1-200; grep -E "norm_|total device" gpurun_out/profile_step_8L.txt | cut -c1-120
