#!/bin/bash
# Profiling batch (1 GPU): launch list of a 4-layer step + ncu --set full of the top kernels.
mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
NCU="ncu --clock-control none"
for t in gemm gemm_wgrad gemm1 norm adam ce paged; do
  python tools/ncu_target.py $t > gpurun_out/plain_$t.log 2>&1 && \
  $NCU --set full --import-source on -k regex:"gemm|norm_|adamw|ce_fwd|paged_attn|swiglu" -s 2 -c 3 -o gpurun_out/prof_$t -f python tools/ncu_target.py $t > gpurun_out/ncu_$t.log 2>&1; echo "ncu_$t rc=$?" >> gpurun_out/rc.txt
done
cat gpurun_out/rc.txt; ls -la gpurun_out/*.ncu-rep 2>/dev/null | head; tail -n 3 gpurun_out/ncu_gemm.log
