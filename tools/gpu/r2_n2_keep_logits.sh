#!/bin/bash
# N=2: lm_head logits kept for the backward (auto) vs recomputed
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "kept_logits" 2>&1 | tail -2
i=0
for cfg in "" "DLA_KEEP_LOGITS=0" ""; do
  i=$((i+1))
  echo "== [$cfg]"
  env $cfg timeout 300 $TR --master-port $((29570+i)) bench.py --gpus 2 --steps 8 --warmup 3 --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['clocks']['sm_mhz'], d['peak_hbm_gb'], d['config']['activation_checkpointing'])"
done
