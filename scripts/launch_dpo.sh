#!/usr/bin/env bash
# dpo phase: one process per GPU (topology from config/accelerate_config.yaml; NPROC overrides).
# usage: scripts/launch_dpo.sh [config.yaml] [extra flags, e.g. --overlay config/ablations/low_lr.yaml --resume latest]
set -euo pipefail
cd "$(dirname "$0")/.."
CONFIG=${1:-config/dpo_config.yaml}
shift || true
export TOKENIZERS_PARALLELISM=false
exec python -m dla_b200.cli.launch ${NPROC:+--nproc $NPROC} --topology config/accelerate_config.yaml \
  -m dla_b200.cli.train_dpo --config "$CONFIG" "$@"
