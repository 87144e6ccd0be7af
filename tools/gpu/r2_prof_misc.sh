#!/bin/bash
# ncu captures of the non-GEMM hot kernels the round-1 review asked for (one GPU)
mkdir -p gpurun_out
for spec in "lmhead:gemm_kernel" "rmsnorm:rmsnorm" "rope:rope_kernel" "adamw:adamw_kernel" "swiglu:swiglu"; do
  w=${spec%%:*}; k=${spec##*:}
  python tools/profile_target.py $w > gpurun_out/plain_$w.log 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 2 -f -o gpurun_out/prof_$w python tools/profile_target.py $w > gpurun_out/ncu_$w.log 2>&1
  echo "ncu $w rc=$?"
done
