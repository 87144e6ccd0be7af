#!/bin/bash
mkdir -p gpurun_out
FILE=tests/test_gpu_kernels.py TESTS="test_top_p_sampler_kernel" TMO=300 bash tools/gpu/kernel_sweep.sh
grep -E "^E  " gpurun_out/test_top_p_sampler_kernel.log | head -5
for w in gemm attn_fwd attn_bwd; do
  python tools/profile_target.py $w > gpurun_out/plain_$w.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:${w}_kernel -s 1 -c 1 -f -o gpurun_out/prof_$w python tools/profile_target.py $w > gpurun_out/ncu_$w.log 2>&1
  echo "ncu $w rc=$?"
done
timeout 600 python tools/profile_step.py > gpurun_out/step_breakdown.log 2>&1; tail -22 gpurun_out/step_breakdown.log | cut -c1-150
