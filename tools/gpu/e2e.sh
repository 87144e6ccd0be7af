#!/bin/bash
mkdir -p gpurun_out
FILE=tests/test_gpu_model.py TESTS="test_model_forward_backward_matches_reference test_dpo_updates_reduce_loss" TMO=300 bash tools/gpu/kernel_sweep.sh
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 3 --warmup 2 > gpurun_out/bench_8b.log 2>&1; echo "bench8b rc=$?"; tail -2 gpurun_out/bench_8b.log | cut -c1-1200
for w in attn_fwd gemm; do
  python tools/profile_target.py $w > gpurun_out/plain_$w.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:${w}_kernel -s 1 -c 1 -f -o gpurun_out/prof_$w python tools/profile_target.py $w > gpurun_out/ncu_$w.log 2>&1
  echo "ncu $w rc=$?"
done
