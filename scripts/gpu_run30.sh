#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/profile_step.py --layers 32 --out gpurun_out/profile_step_32L.txt > gpurun_out/profile_step_32L.log 2>&1; echo "rc=$?"; head -40 gpurun_out/profile_step_32L.txt | cut -c1-170
