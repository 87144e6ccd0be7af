#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest "$@" -x -q -p no:cacheprovider > gpurun_out/one_test.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/one_test.log | cut -c1-250
