#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multigpu.py -x -q -p no:cacheprovider > gpurun_out/multigpu_tests.log 2>&1; echo "multigpu tests rc=$?"; tail -3 gpurun_out/multigpu_tests.log | cut -c1-200
bash tools/gpu/scale_own.sh 2
