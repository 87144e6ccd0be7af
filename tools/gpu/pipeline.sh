#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_pipeline.py -x -q -p no:cacheprovider > gpurun_out/pipeline.log 2>&1; echo "pipeline rc=$?"; tail -25 gpurun_out/pipeline.log
