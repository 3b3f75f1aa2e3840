#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rc.txt
timeout 600 python bench.py --steps 8 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?" >> gpurun_out/rc.txt
timeout 600 python bin/llmctl bench e2e --gen-length 128 --out gpurun_out/bench_e2e_graph.json > gpurun_out/bench_e2e_graph.log 2>&1; echo "e2e rc=$?" >> gpurun_out/rc.txt
SANITIZE_OUT=gpurun_out/sanitize_memcheck.log timeout 600 bash scripts/sanitize.sh memcheck "rmsnorm or rope or swiglu or sumsq or skinny or kv_append or test_flash_attention_fwd" > gpurun_out/sanitize_summary.log 2>&1; echo "sanitize rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 2 gpurun_out/smoke.log | cut -c1-200; tail -n 1 gpurun_out/bench_n1.log | cut -c1-1500; tail -n 18 gpurun_out/bench_e2e_graph.log | cut -c1-200; cat gpurun_out/sanitize_summary.log | tail -5
