#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/kernels.jsonl
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -p no:cacheprovider -k "attention" > gpurun_out/attn2_tests.log 2>&1; echo "attention tests rc=$?"; tail -6 gpurun_out/attn2_tests.log | cut -c1-300
timeout 300 python tools/bench_kernels.py attn > gpurun_out/attn2_bench.log 2>&1; echo "bench rc=$?"; cut -c1-420 gpurun_out/attn2_bench.log | tail -4
