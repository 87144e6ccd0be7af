#!/bin/bash
# 8-GPU round: headline bench (own), Llama-3-70B DPO sizing run, RLHF step with three sharded 8B models
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519"
timeout 600 $TR bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/n8_own.log 2>&1; echo "own 8B N=8 rc=$?"; grep '"metric"' gpurun_out/n8_own.log | cut -c1-400
for comm in auto nccl; do
  DLA_BENCH_NO_RECOMPUTE=0 timeout 600 $TR bench.py --gpus 8 --model llama3-70b --state-dtype bf16 --micro-batch 2 --resident 0 --comm $comm --steps 3 --warmup 3 > gpurun_out/n8_70b_$comm.log 2>&1
  rc=$?; echo "own 70B N=8 comm=$comm rc=$rc"; grep '"metric"' gpurun_out/n8_70b_$comm.log | cut -c1-1200; tail -4 gpurun_out/n8_70b_$comm.log | grep -v metric | cut -c1-300
  [ $rc = 0 ] && break
done
timeout 900 $TR tools/bench_configs.py ppo --model llama3-8b > gpurun_out/n8_ppo.log 2>&1; echo "ppo N=8 rc=$?"; tail -2 gpurun_out/n8_ppo.log | cut -c1-500
nvidia-smi --query-gpu=memory.used --format=csv | head -3
