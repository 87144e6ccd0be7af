#!/bin/bash
# own arm only at N GPUs (headline config)
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/own_n$N.log 2>&1; echo "own N=$N rc=$?"
grep '"metric"' gpurun_out/own_n$N.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['value'], r['ms_per_step'], 'e2e', r['e2e']['value'], 'loss', r.get('loss'), r['clocks'], 'launches', r['gpu_launches'], r['config']['parallelism'], 'peak', r.get('peak_hbm_gb'))"
grep -m3 "Error\|error" gpurun_out/own_n$N.log | cut -c1-300
