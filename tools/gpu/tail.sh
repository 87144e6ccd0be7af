#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/kernels.jsonl
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py -x -q -p no:cacheprovider > gpurun_out/tail_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/tail_tests.log | cut -c1-250
timeout 600 python tools/bench_kernels.py gemm > gpurun_out/tail_gemm.log 2>&1; echo "gemm bench rc=$?"; python - <<'PY'
import json
for l in open("gpurun_out/tail_gemm.log"):
    if l.startswith("{"):
        r=json.loads(l); print(r["tag"], r["M"],r["N"],r["K"], "fwd %.0f (no tail split %.0f) | dgrad %.0f (%.0f) | wgrad %.0f (%.0f) | cublas %.0f"%(r["tflops_2cta"], r["tflops_2cta_no_tail_split"], r["dgrad_tflops_2cta"], r["dgrad_tflops_2cta_no_tail_split"], r["wgrad_tflops_2cta"], r["wgrad_tflops_2cta_no_tail_split"], r["cublas_tflops"]))
PY
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/iter_bench1.log 2>&1; echo "bench.py rc=$?"; grep '"metric"' gpurun_out/iter_bench1.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['value'], r['ms_per_step'], 'e2e', r['e2e']['value'], 'loss', r.get('loss'), r['clocks'], 'launches', r['gpu_launches'])"
