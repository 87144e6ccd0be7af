#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/kernels.jsonl
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -x -q -p no:cacheprovider > gpurun_out/wide_tests.log 2>&1; echo "kernel+model tests rc=$?"; tail -4 gpurun_out/wide_tests.log | cut -c1-300
timeout 300 python tools/bench_kernels.py attn elem > gpurun_out/wide_bench.log 2>&1; echo "bench rc=$?"; cut -c1-330 gpurun_out/wide_bench.log | tail -6
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/wide_bench1.log 2>&1; echo "bench.py rc=$?"; grep '"metric"' gpurun_out/wide_bench1.log | cut -c1-260
