#!/bin/bash
# Llama-3-70B DPO sizing run on 8 GPUs (streamed ZeRO-3, bf16 Adam moments); falls back to NCCL streaming if the fused path fails
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
for comm in auto nccl; do
  DLA_BENCH_NO_RECOMPUTE=0 timeout 600 $TR bench.py --gpus 8 --model llama3-70b --state-dtype bf16 --micro-batch 2 --resident 0 --comm $comm --steps 3 --warmup 3 > gpurun_out/n8_70b_$comm.log 2>&1
  rc=$?; echo "own 70B N=8 comm=$comm rc=$rc"; grep '"metric"' gpurun_out/n8_70b_$comm.log | cut -c1-1500; grep -m2 "OutOfMemoryError\|Error:" gpurun_out/n8_70b_$comm.log | cut -c1-400
  [ $rc = 0 ] && break
done
