#!/bin/bash
mkdir -p gpurun_out
python tools/profile_target.py gemm_wgrad > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm2_kernel -s 1 -c 1 -f -o gpurun_out/prof_wgrad python tools/profile_target.py gemm_wgrad > gpurun_out/ncu_wgrad.log 2>&1; echo "ncu wgrad rc=$?"
python tools/profile_target.py gemm > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm2_kernel -s 1 -c 1 -f -o gpurun_out/prof_gemm2 python tools/profile_target.py gemm > gpurun_out/ncu_gemm2.log 2>&1; echo "ncu fwd rc=$?"
