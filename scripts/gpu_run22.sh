#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 120 python tools/ncu_skinny.py > gpurun_out/ncu_skinny_plain.log 2>&1; echo "plain rc=$?" >> gpurun_out/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_skinny --launch-skip 4 -c 1 -o gpurun_out/skinny_full -f python tools/ncu_skinny.py > gpurun_out/ncu_skinny.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 3 gpurun_out/ncu_skinny.log
