#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
timeout 300 python tools/profile_decode.py gpt-7b eager > gpurun_out/pd_eager.log 2>&1; echo "eager rc=$?" >> gpurun_out/rc.txt
timeout 300 python tools/profile_decode.py gpt-7b graph > gpurun_out/pd_graph.log 2>&1; echo "graph rc=$?" >> gpurun_out/rc.txt
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/decode_launches.csv python tools/profile_decode.py gpt-7b eager > gpurun_out/pd_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/rc.txt
cat gpurun_out/rc.txt; grep round gpurun_out/pd_eager.log gpurun_out/pd_graph.log | cut -c1-400
