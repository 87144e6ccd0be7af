#!/bin/bash
# usage: gpu_scale.sh N  -> own (auto comm) and reference arm at N GPUs
N=${1:-8}
mkdir -p gpurun_out
run() { # impl
  if [ "$N" = 1 ]; then python bench.py --impl $1 --gpus 1 --steps 5 --warmup 3
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl $1 --gpus $N --steps 5 --warmup 3; fi
}
timeout 900 bash -c "$(declare -f run); N=$N; run own" > gpurun_out/scale_own_$N.log 2>&1; echo "own N=$N rc=$?"; grep '"metric"' gpurun_out/scale_own_$N.log | cut -c1-260; tail -3 gpurun_out/scale_own_$N.log | grep -v metric | cut -c1-300
timeout 1200 bash -c "$(declare -f run); N=$N; run reference" > gpurun_out/scale_ref_$N.log 2>&1; echo "ref N=$N rc=$?"; grep '"metric"\|unavailable' gpurun_out/scale_ref_$N.log | cut -c1-260; tail -3 gpurun_out/scale_ref_$N.log | grep -v metric | cut -c1-300
