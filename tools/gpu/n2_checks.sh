#!/bin/bash
# 2-GPU sanity of the streamed ZeRO-3 engine with bf16 Adam moments (the 70B mode, on a 1B model) and the RLHF step
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -p no:cacheprovider -k "adamw" > gpurun_out/adamw_tests.log 2>&1; echo "adamw tests rc=$?"; tail -3 gpurun_out/adamw_tests.log
for st in fp32 bf16; do
  timeout 600 $TR bench.py --gpus 2 --model llama-1b --resident 0 --state-dtype $st --steps 3 --warmup 3 > gpurun_out/stream_$st.log 2>&1; echo "streamed $st rc=$?"; grep '"metric"' gpurun_out/stream_$st.log | cut -c1-330; tail -2 gpurun_out/stream_$st.log | grep -v metric | cut -c1-300
done
timeout 900 $TR tools/bench_configs.py ppo --model llama-1b > gpurun_out/ppo_n2.log 2>&1; echo "ppo rc=$?"; tail -2 gpurun_out/ppo_n2.log | cut -c1-500
