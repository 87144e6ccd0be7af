#!/bin/bash
# full single-GPU suite + ncu evidence for the new kernels
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q -p no:cacheprovider --ignore=tests/test_multigpu.py > gpurun_out/full_tests.log 2>&1; echo "full gpu tests rc=$?"; tail -4 gpurun_out/full_tests.log | cut -c1-200
python tools/profile_target.py attn_fwd > /dev/null 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_fwd2_kernel -s 1 -c 1 -f -o gpurun_out/prof_attn_fwd2 python tools/profile_target.py attn_fwd > gpurun_out/ncu_attn_fwd2.log 2>&1; echo "ncu attn rc=$?"
