#!/bin/bash
mkdir -p gpurun_out
timeout 100 python tools/trace_flash.py > gpurun_out/trace_flash.log 2>&1; echo "rc=$?"; cat gpurun_out/trace_flash.log | cut -c1-200 | head -120
