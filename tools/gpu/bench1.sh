#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_8b.log 2>&1; echo "bench8b rc=$?"; tail -1 gpurun_out/bench_8b.log | cut -c1-1300
tail -5 gpurun_out/bench_8b.log | grep -v metric | cut -c1-300
