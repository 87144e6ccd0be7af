#!/bin/bash
# A/B of the side-stream optimizer overlap at N GPUs (default 1) + engine/pipeline tests
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = 1 ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"; fi
if [ "$N" = 1 ]; then
  timeout 1200 python -m pytest tests/test_gpu_model.py tests/test_gpu_pipeline.py -x -q -p no:cacheprovider > gpurun_out/overlap_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/overlap_tests.log | cut -c1-200
else
  timeout 1200 python -m pytest tests/test_multigpu.py -x -q -p no:cacheprovider > gpurun_out/overlap_tests.log 2>&1; echo "multigpu tests rc=$?"; tail -3 gpurun_out/overlap_tests.log | cut -c1-200
fi
for ov in 1 0; do
  DLA_OVERLAP_OPT=$ov timeout 600 $RUN bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/overlap_${ov}_$N.log 2>&1; echo "overlap=$ov N=$N rc=$?"; grep '"metric"' gpurun_out/overlap_${ov}_$N.log | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['value'], r['ms_per_step'], 'e2e', r['e2e']['value'], 'loss', r.get('loss'), r['clocks'])"
done
