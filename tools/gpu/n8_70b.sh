#!/bin/bash
# Llama-3-70B DPO on 8 GPUs (streamed ZeRO-3, bf16 Adam moments), micro-batch MB pairs per GPU
MB=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
DLA_BENCH_NO_RECOMPUTE=0 timeout 600 $TR bench.py --gpus 8 --model llama3-70b --state-dtype bf16 --micro-batch $MB --resident 0 --steps 3 --warmup 3 > gpurun_out/n8_70b_mb$MB.log 2>&1
rc=$?; echo "own 70B N=8 mb=$MB rc=$rc"; grep '"metric"' gpurun_out/n8_70b_mb$MB.log | cut -c1-1500; grep -m2 "OutOfMemoryError\|Error:" gpurun_out/n8_70b_mb$MB.log | cut -c1-400
