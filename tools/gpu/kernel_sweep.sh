#!/bin/bash
# Runs each GPU kernel test in its own process under a timeout so one hang cannot hide the rest.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
TESTS=${TESTS:-"test_rmsnorm_fwd_bwd test_layernorm_fwd_bwd test_rope_qkv test_swiglu_gelu test_embedding test_adamw_and_norm test_losses test_gemm_nt test_gemm_layout_variants test_linear_autograd test_linear_logprob"}
FILE=${FILE:-tests/test_gpu_kernels.py}
: > gpurun_out/sweep_summary.txt
for t in $TESTS; do
  timeout ${TMO:-240} python -m pytest $FILE -k "$t" -q --no-header -x -p no:cacheprovider > gpurun_out/$t.log 2>&1
  rc=$?
  echo "$t rc=$rc $(tail -1 gpurun_out/$t.log)" >> gpurun_out/sweep_summary.txt
done
cat gpurun_out/sweep_summary.txt
