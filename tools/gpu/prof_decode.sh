#!/bin/bash
mkdir -p gpurun_out
for w in skinny:skinny_kernel decode_attn:decode_attn_kernel; do
  t=${w%%:*}; k=${w##*:}
  python tools/profile_target.py $t > gpurun_out/plain_$t.log 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o gpurun_out/prof_$t python tools/profile_target.py $t > gpurun_out/ncu_$t.log 2>&1
  echo "ncu $t rc=$?"
done
