#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/kernels.jsonl
timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -p no:cacheprovider -k "early_s" > gpurun_out/attn3_tests.log 2>&1; echo "fwd3 tests rc=$?"; tail -6 gpurun_out/attn3_tests.log | cut -c1-300
timeout 300 python tools/bench_kernels.py attn > gpurun_out/attn3_bench.log 2>&1; echo "bench rc=$?"; cut -c1-460 gpurun_out/attn3_bench.log | tail -3
