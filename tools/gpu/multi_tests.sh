#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_multigpu.py -x -q -p no:cacheprovider > gpurun_out/multigpu_tests.log 2>&1; echo "multigpu tests rc=$?"; tail -12 gpurun_out/multigpu_tests.log | cut -c1-250
