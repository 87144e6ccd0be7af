#!/bin/bash
mkdir -p gpurun_out
FILE=tests/test_gpu_kernels.py TESTS="test_gemm_nt test_gemm_layout_variants test_linear_autograd test_linear_logprob" TMO=300 bash tools/gpu/kernel_sweep.sh
grep -E "^E  |Error" gpurun_out/test_gemm_*.log | head -10
FILE=tests/test_zz_no_alloc.py TESTS="test_noalloc" TMO=300 bash tools/gpu/kernel_sweep.sh
timeout 300 python tools/bench_kernels.py gemm 2>&1 | tail -8 | cut -c1-420
timeout 900 python bench.py --steps 3 --warmup 2 > gpurun_out/bench_8b.log 2>&1; echo "bench8b rc=$?"; tail -1 gpurun_out/bench_8b.log | cut -c1-400
