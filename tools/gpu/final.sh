#!/bin/bash
# End-of-round validation on a 2-GPU box: the whole GPU suite (multi-GPU tests at world 2), smoke(), a short N=2 bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/final_gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_gpu_tests.log
tail -4 gpurun_out/final_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/final_smoke.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/final_bench_n2.json 2> gpurun_out/final_bench_n2.err; echo "bench rc=$?"
tail -1 gpurun_out/final_bench_n2.json | cut -c1-400
