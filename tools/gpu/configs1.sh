#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
for impl in own reference; do
  timeout 600 python tools/bench_configs.py latency --impl $impl > gpurun_out/lat_$impl.log 2>&1; echo "latency $impl rc=$?"; tail -2 gpurun_out/lat_$impl.log | cut -c1-250
  timeout 900 python tools/bench_configs.py sft --impl $impl > gpurun_out/sft_$impl.log 2>&1; echo "sft $impl rc=$?"; tail -2 gpurun_out/sft_$impl.log | cut -c1-300
done
