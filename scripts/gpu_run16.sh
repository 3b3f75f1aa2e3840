#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 120 python tools/ncu_attention.py > gpurun_out/ncu_attn_plain.log 2>&1; echo "plain rc=$?" >> gpurun_out/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"flash_fwd2|flash_bwd2_kernel" --launch-skip 3 -c 3 -o gpurun_out/attn_full -f python tools/ncu_attention.py > gpurun_out/ncu_attn.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; tail -n 5 gpurun_out/ncu_attn.log
