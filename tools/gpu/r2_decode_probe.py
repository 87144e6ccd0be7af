"""Per-kernel times of one decode token (torch.profiler over a few graph replays; prefill is averaged in)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
from dla_b200.models.factory import load_causal_lm  # noqa: E402
from dla_b200.ops import functional as F  # noqa: E402
from dla_b200.parallel.pg import init_distributed  # noqa: E402
from dla_b200.rollout import Generator  # noqa: E402

ctx = init_distributed()
dev = ctx.device
r = lambda *s: torch.randn(*s, device=dev, dtype=torch.bfloat16)


def t(fn, n=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


with torch.no_grad():
    b = load_causal_lm("llama3-8b", ctx, trainable=False, gradient_checkpointing=False, name="gen")
    ids = torch.randint(4, b.config.vocab_size, (8, 256), device=dev)
    gen = Generator(b.model)
    run = lambda n: gen.generate(ids, torch.ones_like(ids), max_new_tokens=n, do_sample=True, temperature=0.7, top_p=0.9, seed=1,
                                 eos_token_id=-1, pad_token_id=0)
    run(8)
    torch.cuda.synchronize()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        run(17)
        torch.cuda.synchronize()
    rows = sorted(((e.key[:70], e.device_time_total / 16, e.count / 16) for e in prof.key_averages() if e.device_time_total > 0),
                  key=lambda z: -z[1])
    tot = sum(z[1] for z in rows)
    print(f"sum of kernel time per decode token (prefill included in the average, 16 tokens): {tot / 1e3:.2f} ms")
    for k, us, c in rows[:16]:
        print(f"  {us:8.1f} us/token  x{c:6.1f}  {k}")
