#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider --ignore=tests/test_multigpu.py > gpurun_out/full_tests.log 2>&1; echo "full gpu tests rc=$?"; tail -3 gpurun_out/full_tests.log | cut -c1-200
timeout 900 python tools/bench_configs.py decode --impl own > gpurun_out/decode_own.log 2>&1; echo "decode own rc=$?"; tail -2 gpurun_out/decode_own.log | cut -c1-330
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/iter_bench1.log 2>&1; echo "bench.py rc=$?"; grep '"metric"' gpurun_out/iter_bench1.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['value'], r['ms_per_step'], 'e2e', r['e2e']['value'], 'loss', r.get('loss'), r['clocks'], 'launches', r['gpu_launches'])"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
