#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kernels.py -x -q -p no:cacheprovider -k "scalar_head or rl_objective or distill" > gpurun_out/newk_tests.log 2>&1; echo "new kernel tests rc=$?"; tail -12 gpurun_out/newk_tests.log
timeout 1500 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_model.py -x -q -p no:cacheprovider > gpurun_out/pipeline_tests.log 2>&1; echo "pipeline rc=$?"; tail -12 gpurun_out/pipeline_tests.log
