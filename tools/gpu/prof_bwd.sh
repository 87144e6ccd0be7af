#!/bin/bash
mkdir -p gpurun_out
python tools/profile_target.py attn_bwd > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_bwd_kernel -s 1 -c 1 -f -o gpurun_out/prof_attn_bwd2 python tools/profile_target.py attn_bwd > gpurun_out/ncu_attn_bwd2.log 2>&1; echo "ncu bwd rc=$?"
