"""Developer A/B timing of tuning variants INSIDE one process, interleaved round by round, because a
B200 under its power cap drifts by 10-20 % over a minute and back-to-back processes do not compare.
  python tests/cuda/ab_bench.py fwd|bwd [variant ...]      ('' = the product library is always included)
Env: AB_P (dropout, default 0.1), AB_CAUSAL (1), AB_ROUNDS (6), AB_ITERS (10), AB_B/AB_S/AB_H/AB_D."""
import os, sys, statistics
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

which = sys.argv[1]
variants = [""] + sys.argv[2:]
p = float(os.environ.get("AB_P", 0.1)); causal = bool(int(os.environ.get("AB_CAUSAL", 1)))
rounds = int(os.environ.get("AB_ROUNDS", 6)); iters = int(os.environ.get("AB_ITERS", 10))
b, s, h, d = (int(os.environ.get(k, v)) for k, v in (("AB_B", 16), ("AB_S", 4096), ("AB_H", 32), ("AB_D", 128)))
hk = int(os.environ.get("AB_HK", h))

libs = {}
for v in variants:
    api._lib = None
    api.LIB_PATH = os.path.join(os.path.dirname(api.__file__), "lib", "libflash_attn" + ("_" + v if v else "") + ".so")
    libs[v] = api.lib()

q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16); v_ = torch.randn_like(k); do = torch.randn_like(q)
out = torch.empty_like(q)
kw = dict(p_dropout=p, is_causal=causal, seed=1234, offset=0)
api._lib = libs[""]
_, lse, _, _ = api.flash_attn_fwd(q, k, v_, out=out, **kw)
scratch = (torch.empty((b, h, s), dtype=torch.float32, device="cuda"),
           torch.empty((b, s, h, api.round_up(d, 32)), dtype=torch.float32, device="cuda"))
grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v_))

def run():
    if which == "fwd":
        api.flash_attn_fwd(q, k, v_, out=out, **kw)
    else:
        api.flash_attn_bwd(do, q, k, v_, out, lse, scratch=scratch, grads=grads, **kw)

fl = 4.0 * b * h * s * s * d * (0.5 if causal else 1.0) * (1.0 if which == "fwd" else 2.5)
ms = {v: [] for v in variants}
for r in range(rounds + 1):
    for v in variants:
        api._lib = libs[v]
        run(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            run()
        e1.record(); torch.cuda.synchronize()
        if r:                     # round 0 warms up
            ms[v].append(e0.elapsed_time(e1) / iters)
print(f"{which} b{b} s{s} h{h}/{hk} d{d} causal={causal} p={p}: {rounds} interleaved rounds x {iters} launches")
for v in variants:
    m = statistics.median(ms[v])
    print(f"  {v or '(product)':12s} median {m:.3f} ms {fl / m / 1e9:7.1f} TFLOP/s   best {min(ms[v]):.3f} ms {fl / min(ms[v]) / 1e9:7.1f}   "
          + " ".join(f"{x:.3f}" for x in ms[v]), flush=True)
