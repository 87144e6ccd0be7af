#!/bin/bash
mkdir -p gpurun_out
FILE=tests/test_gpu_kernels.py TESTS="test_attention_fwd_bwd test_attention_separate_tensors" TMO=300 bash tools/gpu/kernel_sweep.sh
FILE=tests/test_gpu_model.py TESTS="test_model_forward_backward_matches_reference" TMO=300 bash tools/gpu/kernel_sweep.sh
timeout 300 python tools/bench_kernels.py attn 2>&1 | tail -3
timeout 900 python bench.py --steps 3 --warmup 2 > gpurun_out/bench_8b.log 2>&1; echo "bench8b rc=$?"; tail -1 gpurun_out/bench_8b.log | cut -c1-400
