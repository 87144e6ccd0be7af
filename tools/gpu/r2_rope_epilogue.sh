#!/bin/bash
# RoPE in the QKV epilogue + out-of-place inverse rotation: tests, same-box A/B of the headline step
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py -x -q -k "rope or gemm or linear" > gpurun_out/r2_rope_tests.log 2>&1; echo "kernels rc=$?" >> gpurun_out/r2_rope_tests.log
python -m pytest tests/test_gpu_model.py tests/test_cli_cpu.py -x -q -m gpu >> gpurun_out/r2_rope_tests.log 2>&1; echo "model rc=$?" >> gpurun_out/r2_rope_tests.log
grep -E "passed|failed|rc=|Error|error" gpurun_out/r2_rope_tests.log | tail -8
DLA_FUSE_ROPE=0 python bench.py --steps 8 --warmup 3 --no-e2e > gpurun_out/r2_n1_rope_separate.json 2> gpurun_out/r2_n1_rope_separate.err
python bench.py --steps 8 --warmup 3 --no-e2e > gpurun_out/r2_n1_rope_epilogue.json 2> gpurun_out/r2_n1_rope_epilogue.err
python - <<'PY'
import json
for f in ("r2_n1_rope_separate", "r2_n1_rope_epilogue"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["ms_per_step"], 2), d["clocks"]["sm_mhz"], d["loss"])
    except Exception as e:
        print(f, "ERR", e, open(f"gpurun_out/{f}.err").read()[-600:])
PY
