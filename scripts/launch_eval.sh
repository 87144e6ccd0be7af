#!/usr/bin/env bash
# evaluation: alignment proxies, then the prefill latency grid (single process, like the reference)
set -euo pipefail
cd "$(dirname "$0")/.."
CONFIG=${1:-config/eval_config.yaml}
export TOKENIZERS_PARALLELISM=false
python -m dla_b200.cli.eval_alignment --config "$CONFIG"
python -m dla_b200.cli.eval_latency --config "$CONFIG"
