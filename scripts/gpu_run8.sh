#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 120 --timeout-method=thread > gpurun_out/t_noalloc_alone.log 2>&1; echo "noalloc_alone rc=$?" >> gpurun_out/rc.txt
LLMCTL_DEBUG=1 timeout 900 python -m pytest tests -m gpu -q -x --timeout 200 --timeout-method=thread > gpurun_out/t_all.log 2>&1; echo "all rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 5 gpurun_out/t_noalloc_alone.log; tail -n 8 gpurun_out/t_all.log
