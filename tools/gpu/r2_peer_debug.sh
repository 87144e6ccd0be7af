#!/bin/bash
# Multi-GPU tests against a library built with the signal-pad sequence assertions (needs >= 2 GPUs)
set -e
DLA_BUILD_DEFINES="-DDLA_PEER_DEBUG" python -m dla_b200.ops.build --force > /dev/null
DLA_BUILD_DEFINES="-DDLA_PEER_DEBUG" timeout 900 python -m pytest tests/test_multigpu.py -x -q 2>&1 | tail -4
