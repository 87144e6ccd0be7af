#!/bin/bash
# bias-in-epilogue GEMM + in-backward grad norm (N=1): tests, same-box A/B of the headline step, step breakdown
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py -x -q -k "gemm or linear or mlp or swiglu or lm_head" > gpurun_out/r2_bias_tests.log 2>&1
echo "kernels rc=$?" >> gpurun_out/r2_bias_tests.log
python -m pytest tests/test_gpu_model.py tests/test_gpu_property.py -x -q >> gpurun_out/r2_bias_tests.log 2>&1
echo "model rc=$?" >> gpurun_out/r2_bias_tests.log
tail -4 gpurun_out/r2_bias_tests.log
DLA_NORM_IN_BACKWARD=0 python bench.py --steps 8 --warmup 3 --no-e2e > gpurun_out/r2_n1_norm_after.json 2> gpurun_out/r2_n1_norm_after.err
python bench.py --steps 8 --warmup 3 > gpurun_out/r2_n1_norm_in_bwd.json 2> gpurun_out/r2_n1_norm_in_bwd.err
python tools/profile_step.py > gpurun_out/r2_step_breakdown.log 2>&1
tail -2 gpurun_out/r2_n1_norm_after.json gpurun_out/r2_n1_norm_in_bwd.json
head -30 gpurun_out/r2_step_breakdown.log
