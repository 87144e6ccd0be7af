#!/bin/bash
# ncu captures of the fused-epilogue GEMMs, the lm_head log-prob GEMM and the all-gather GEMM (one GPU, world = 1)
mkdir -p gpurun_out
for spec in "lmhead:gemm_kernel:1:1" "qkv_rope:gemm2_kernel:1:1" "mlp_epilogues:gemm2_kernel:6:6" "ag_gemm:gemm_kernel:1:1"; do
  IFS=: read w k s c <<< "$spec"
  python tools/profile_target.py $w > gpurun_out/plain_$w.log 2>&1 && \
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:$k -s $s -c $c -f -o gpurun_out/prof_$w python tools/profile_target.py $w > gpurun_out/ncu_$w.log 2>&1
  echo "ncu $w rc=$? $(tail -1 gpurun_out/plain_$w.log)"
done
ls -la gpurun_out/prof_*.ncu-rep
