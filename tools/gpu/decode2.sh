#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/configs.jsonl
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_pipeline.py -x -q -p no:cacheprovider -k "decode or rollout" > gpurun_out/decode_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/decode_tests.log
python - <<'PY'
import torch, sys
sys.path.insert(0, ".")
from dla_b200.ops import functional as F, native as N
dev="cuda"; r=lambda *s: torch.randn(*s, device=dev, dtype=torch.bfloat16)
B,S,Hq,Hk,D=8,320,32,8,128
qkv=r(B,(Hq+2*Hk)*D); kc,vc=r(B,S,Hk,D),r(B,S,Hk,D)
lens=torch.full((B,),300,device=dev,dtype=torch.int32)
cos,sin=F.rope_tables(8192,D,500000.0,torch.device(dev)); o=torch.empty(B,Hq,D,device=dev,dtype=torch.bfloat16)
f=lambda: N.call("dla_decode_attn_fused",N.ptr(qkv),qkv.stride(0),N.ptr(kc),N.ptr(vc),N.ptr(o),N.ptr(lens),N.ptr(cos),N.ptr(sin),B,S,Hq,Hk,D,D,D**-0.5,N.stream_ptr())
for _ in range(5): f()
torch.cuda.synchronize()
a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(100): f()
b.record(); torch.cuda.synchronize()
print("decode_attn_fused us/call (back-to-back):", a.elapsed_time(b)*10)
PY
timeout 900 python tools/bench_configs.py decode --impl own > gpurun_out/decode_own.log 2>&1; echo "decode own rc=$?"; tail -2 gpurun_out/decode_own.log | cut -c1-330
