#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -p no:cacheprovider -k "gemm" > gpurun_out/ab_tests.log 2>&1; echo "gemm tests rc=$?"; tail -2 gpurun_out/ab_tests.log | cut -c1-200
for k in 16384 32768; do
  DLA_GEMM_2CTA_MAX_K=$k timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e > gpurun_out/ab_k_$k.log 2>&1; echo "max_k=$k rc=$?"; grep '"metric"' gpurun_out/ab_k_$k.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['value'], r['ms_per_step'], r['clocks'])"
done
