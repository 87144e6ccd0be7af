#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -x -p no:cacheprovider > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x -p no:cacheprovider > gpurun_out/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/gpu_suite.log
