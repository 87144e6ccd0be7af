#!/bin/bash
# Llama-3-70B DPO on 8 GPUs, same box back to back: fused bucket-ring reduce-scatter (default) vs NCCL reduce-scatter
# (DLA_STREAM_RS=0, the round-1 configuration); the parameter all-gather is fused into the GEMMs in both
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
i=0
for cfg in "DLA_STREAM_RS=1" "DLA_STREAM_RS=0"; do
  i=$((i+1))
  echo "== [$cfg]"
  env $cfg DLA_BENCH_NO_RECOMPUTE=0 timeout 500 $TR --master-port $((29580+i)) bench.py --gpus 8 --model llama3-70b --state-dtype bf16 --micro-batch 4 --resident 0 --steps 3 --warmup 3 --no-e2e 2> gpurun_out/r2_70b_ab_$i.err | tail -1 > gpurun_out/r2_70b_ab_$i.json
  python -c "import sys,json; d=json.loads(open('gpurun_out/r2_70b_ab_$i.json').read()); print(d['ms_per_step'], d['value'], d['clocks'], d['peak_hbm_gb'], d['loss'])" || tail -3 gpurun_out/r2_70b_ab_$i.err
done
