#!/bin/bash
# N=8: in-backward reduce-scatter (default) vs the round-1 order, same box, back to back; per-rank busy clocks
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
i=0
for cfg in "" "DLA_RS_IN_BACKWARD=0"; do
  i=$((i+1))
  echo "== [$cfg]"
  nvidia-smi --query-gpu=index,clocks.sm,power.draw --format=csv,noheader -lms 500 > gpurun_out/r2_n8_ab_smi_$i.csv &
  SMI=$!
  env $cfg timeout 300 $TR --master-port $((29550+i)) bench.py --gpus 8 --steps 6 --warmup 3 --no-e2e 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['clocks'], d['config'].get('comm'))"
  kill $SMI
  python - <<PY
import collections, statistics
rows = collections.defaultdict(list)
for ln in open("gpurun_out/r2_n8_ab_smi_$i.csv"):
    f = [x.strip() for x in ln.split(",")]
    try: rows[int(f[0])].append((float(f[1].split()[0]), float(f[2].split()[0])))
    except Exception: pass
for g, v in sorted(rows.items()):
    busy = [c for c, p in v if p > 600]
    print(f"  gpu{g}: median SM clock under load {statistics.median(busy) if busy else None} MHz, median power {statistics.median([p for c,p in v if p > 600]) if busy else None} W, samples {len(busy)}")
PY
done
