#!/bin/bash
# round-end style validation on one GPU: smoke, full GPU suite, own bench, reference-arm bench
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/final_smoke.log
timeout 2400 python -m pytest tests -m gpu -x -q -p no:cacheprovider > gpurun_out/final_tests.log 2>&1; echo "pytest -m gpu rc=$?"; tail -3 gpurun_out/final_tests.log | cut -c1-200
timeout 900 python bench.py --gpus 1 > gpurun_out/final_bench_own.log 2>&1; echo "bench own rc=$?"; grep '"metric"' gpurun_out/final_bench_own.log | cut -c1-1500
timeout 900 python bench.py --impl reference --gpus 1 > gpurun_out/final_bench_ref.log 2>&1; echo "bench ref rc=$?"; grep '"metric"\|unavailable' gpurun_out/final_bench_ref.log | cut -c1-600
