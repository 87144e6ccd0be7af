#!/bin/bash
mkdir -p gpurun_out
FILE=tests/test_gpu_kernels.py TESTS="test_gemm_2cta_variant" TMO=200 bash tools/gpu/kernel_sweep.sh
grep -E "^E  |Error|rror" gpurun_out/test_gemm_2cta_variant.log | head -8
timeout 400 python tools/bench_kernels.py gemm 2>&1 | tail -8 | python -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l[:200]); continue
    print(d['tag'], 'own', round(d['tflops']), '2cta', round(d['tflops_2cta']), 'cublas', round(d['cublas_tflops']), '| dgrad', round(d['dgrad_tflops']), round(d['dgrad_tflops_2cta']), '| wgrad', round(d['wgrad_tflops']), round(d['wgrad_tflops_2cta']))"
