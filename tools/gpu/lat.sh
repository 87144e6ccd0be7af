#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
timeout 900 python -m pytest tests/test_gpu_pipeline.py -x -q -p no:cacheprovider > gpurun_out/pipe_tests.log 2>&1; echo "pipeline tests rc=$?"; tail -3 gpurun_out/pipe_tests.log | cut -c1-250
timeout 600 python tools/bench_configs.py latency --impl own > gpurun_out/lat_own.log 2>&1; echo "latency own rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/configs.jsonl'):
    r=json.loads(l); print(r['batch_size'],r['seq_length'],round(r['latency_ms'],2),round(r['tokens_per_second']))
PY
