#!/bin/bash
# usage: gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu.py -x -q -p no:cacheprovider > gpurun_out/multigpu_tests.log 2>&1; echo "multigpu tests rc=$?"; tail -15 gpurun_out/multigpu_tests.log
for comm in fused nccl; do
  DLA_BENCH_COMM=$comm timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 2 > gpurun_out/bench_own_${comm}_$N.log 2>&1
  echo "bench own $comm N=$N rc=$?"; tail -1 gpurun_out/bench_own_${comm}_$N.log | cut -c1-900
done
