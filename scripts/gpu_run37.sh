#!/bin/bash
mkdir -p gpurun_out/tuning_results
timeout 150 python bin/llmctl tune full --device cuda --max-iterations 12 --timeout 60 --output-dir gpurun_out/tuning_results > gpurun_out/tune_full.log 2>&1; echo "rc=$?"; tail -n 30 gpurun_out/tune_full.log | cut -c1-150
