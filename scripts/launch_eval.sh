#!/usr/bin/env bash
# Evaluation phase.  Two single-process passes over the same YAML: (1) alignment proxies per model x benchmark
# (results.json + summary.md), (2) the prefill latency grid (latency.json; one CUDA graph per batch x sequence shape).
# usage: scripts/launch_eval.sh [config/eval_config.yaml]
set -euo pipefail
repo_root="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
cd "$repo_root"
cfg="${1:-config/eval_config.yaml}"

export TOKENIZERS_PARALLELISM=false
for module in eval_alignment eval_latency; do
  python -m "dla_b200.cli.${module}" --config "$cfg"
done
