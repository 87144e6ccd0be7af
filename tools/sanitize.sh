#!/usr/bin/env bash
# Race / memory checking of the native kernels on one GPU.  One tool per invocation — running
# several compute-sanitizer tools in one call has left B200 boxes unusable (B200_PROFILING.md).  On the round-1 pool
# compute-sanitizer was closed altogether (gpurun answers rc 86); bounds are then covered by the odd-size cases of
# tests/test_gpu_kernels.py against fp32 PyTorch.
# The cross-GPU signal-pad protocol is outside any sanitizer's view; its checker is a debug build with sequence-number
# assertions (csrc/peer.cu DLA_PEER_ASSERT): DLA_BUILD_DEFINES="-DDLA_PEER_DEBUG" python -m dla_b200.ops.build --force,
# then the multi-GPU tests (tools/gpu/r2_peer_debug.sh).
#   tools/sanitize.sh memcheck|racecheck|synccheck|initcheck [pytest -k expression]
set -euo pipefail
cd "$(dirname "$0")/.."
TOOL=${1:-memcheck}
EXPR=${2:-"rmsnorm or swiglu or rope or losses or adamw"}
exec compute-sanitizer --tool "$TOOL" --error-exitcode 1 python -m pytest tests/test_gpu_kernels.py -x -q -k "$EXPR" -p no:cacheprovider
