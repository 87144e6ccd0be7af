#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
timeout 600 python tools/bench_configs.py sft --impl own > gpurun_out/sft_own.log 2>&1; echo "sft own rc=$?"; tail -1 gpurun_out/sft_own.log | cut -c1-300
timeout 600 python tools/bench_configs.py latency --impl own > gpurun_out/lat_own.log 2>&1; echo "latency own rc=$?"; tail -3 gpurun_out/lat_own.log | cut -c1-250
