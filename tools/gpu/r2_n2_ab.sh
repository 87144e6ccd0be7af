#!/bin/bash
# A/B of the multi-GPU gradient-exchange schedules at N=2 (cheap): in-backward reduce-scatter (default), after-backward
# (round-1 order), peer loads/stores instead of multimem, and the NCCL engine
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
i=0
for cfg in "" "DLA_RS_IN_BACKWARD=0" "DLA_NO_MULTICAST=1" "DLA_BENCH_COMM=nccl"; do
  i=$((i+1))
  echo "== [$cfg]"
  env $cfg timeout 300 $TR --master-port $((29540+i)) bench.py --gpus 2 --steps 8 --warmup 3 --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['clocks'], d['config'].get('comm'))"
done
