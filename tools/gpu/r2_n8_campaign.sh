#!/bin/bash
# Round-2 N=8 measurement campaign (one 8-GPU box): headline bench, RLHF configs, SFT-4k own + reference, Llama-3-70B.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
run() { echo "== $1"; shift; "$@"; echo "rc=$?"; }
run "bench own N=8" bash -c "timeout 400 $TR --master-port 29531 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/r2c_own_n8.json 2> gpurun_out/r2c_own_n8.err; tail -1 gpurun_out/r2c_own_n8.json | cut -c1-600"
run "ppo reinforce N=8" bash -c "timeout 400 $TR --master-port 29532 tools/bench_configs.py ppo --algo reinforce --steps 3 --warmup 2 2>&1 | grep '\"config\"' | cut -c1-700"
run "ppo clipped N=8" bash -c "timeout 400 $TR --master-port 29533 tools/bench_configs.py ppo --algo ppo --steps 3 --warmup 2 2>&1 | grep '\"config\"' | cut -c1-700"
run "sft 4k own N=8" bash -c "timeout 400 $TR --master-port 29534 tools/bench_configs.py sft --steps 4 --warmup 2 2>&1 | grep '\"config\"' | cut -c1-500"
run "sft 4k reference N=8" bash -c "timeout 500 $TR --master-port 29535 tools/bench_configs.py sft --impl reference --steps 4 --warmup 2 2>&1 | grep '\"config\"' | cut -c1-500"
run "70B own fused N=8" bash -c "DLA_BENCH_NO_RECOMPUTE=0 timeout 500 $TR --master-port 29536 bench.py --gpus 8 --model llama3-70b --state-dtype bf16 --micro-batch 4 --resident 0 --steps 3 --warmup 3 --no-e2e > gpurun_out/r2c_70b_fused.json 2> gpurun_out/r2c_70b_fused.err; tail -1 gpurun_out/r2c_70b_fused.json | cut -c1-700; grep -m2 'Error' gpurun_out/r2c_70b_fused.err | cut -c1-300"
