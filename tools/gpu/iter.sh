#!/bin/bash
# quick iteration: kernel + model tests, elementwise/attention bench, headline bench at N=1
mkdir -p gpurun_out; rm -f gpurun_out/kernels.jsonl
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_pipeline.py -x -q -p no:cacheprovider > gpurun_out/iter_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/iter_tests.log | cut -c1-300
timeout 300 python tools/bench_kernels.py elem ${1:-} > gpurun_out/iter_bench.log 2>&1; echo "bench rc=$?"; cut -c1-330 gpurun_out/iter_bench.log | tail -5
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/iter_bench1.log 2>&1; echo "bench.py rc=$?"; grep '"metric"' gpurun_out/iter_bench1.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['value'], r['ms_per_step'], 'e2e', r['e2e']['value'], 'loss', r.get('loss'), r['clocks'], 'launches', r['gpu_launches'])"
