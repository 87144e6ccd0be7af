#!/bin/bash
# usage: gpu_decode.sh [ref]  -> decode-path tests, skinny GEMM bench, decode ms/token (own; reference arm when "ref")
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl gpurun_out/kernels.jsonl
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_pipeline.py -x -q -p no:cacheprovider -k "skinny or rollout or sampler or gemm_nt or decode" > gpurun_out/decode_tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/decode_tests.log
timeout 600 python tools/bench_kernels.py skinny > gpurun_out/skinny.log 2>&1; echo "skinny rc=$?"; python - <<'PY'
import json
for l in open("gpurun_out/skinny.log"):
    if l.startswith("{"):
        r = json.loads(l); print(r["tag"], r["M"], "ms %.4f  %.0f GB/s  tile %.4f  cublas %.4f" % (r["ms"], r["gbs"], r["tile_kernel_ms"], r["cublas_ms"]))
PY
for impl in own $1; do
  [ "$impl" = ref ] && impl=reference
  timeout 900 python tools/bench_configs.py decode --impl $impl > gpurun_out/decode_$impl.log 2>&1; echo "decode $impl rc=$?"; tail -3 gpurun_out/decode_$impl.log | cut -c1-330
done
