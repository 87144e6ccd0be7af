#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/kernels.jsonl gpurun_out/configs.jsonl
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_pipeline.py -x -q -p no:cacheprovider > gpurun_out/splitk_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/splitk_tests.log | cut -c1-250
timeout 300 python tools/bench_kernels.py smallm > gpurun_out/smallm.log 2>&1; echo "smallm rc=$?"; python - <<'PY'
import json
for l in open("gpurun_out/smallm.log"):
    if l.startswith("{"):
        r=json.loads(l); print(r["tag"], r["M"], "split %.4f  nosplit %.4f  cublas %.4f ms  (%.0f TFLOP/s, W %.0f GB/s)"%(r["ms_splitk"], r["ms_no_split"], r["cublas_ms"], r["tflops"], r["weight_gbs"]))
PY
timeout 600 python tools/bench_configs.py latency --impl own > gpurun_out/lat_own.log 2>&1; echo "latency own rc=$?"; python - <<'PY'
import json
for l in open('gpurun_out/configs.jsonl'):
    r=json.loads(l); print(r['batch_size'],r['seq_length'],round(r['latency_ms'],2),round(r['tokens_per_second']))
PY
