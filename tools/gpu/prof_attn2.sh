#!/bin/bash
mkdir -p gpurun_out
python tools/profile_target.py attn_fwd > gpurun_out/plain_attn_fwd2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fwd2_kernel -s 1 -c 1 -f -o gpurun_out/prof_attn_fwd2 python tools/profile_target.py attn_fwd > gpurun_out/ncu_attn_fwd2.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/prof_attn_fwd2.ncu-rep
