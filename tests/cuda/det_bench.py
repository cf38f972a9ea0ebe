"""Deterministic (num_splits == 1) against arrival-order backward, device-timed through the C ABI."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

def run(b, s, h, hk, d, causal, p, mask=False, iters=10):
    q, do = (torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16) for _ in range(2))
    k, v = (torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16) for _ in range(2))
    m = torch.randn(1, 1, s, s, device="cuda", dtype=torch.bfloat16) if mask else None
    kw = dict(p_dropout=p, is_causal=causal, seed=5, offset=1, attn_mask=m)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, **kw)
    sq_r, d_r = api.round_up(s, 128), api.round_up(d, 32)
    scratch = (torch.empty((b, h, sq_r), dtype=torch.float32, device="cuda"), torch.empty((b, sq_r, h, d_r), dtype=torch.float32, device="cuda"))
    grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    res = []
    for det in (False, True):
        f = lambda: api.flash_attn_bwd(do, q, k, v, out, lse, scratch=scratch, grads=grads, deterministic=det, **kw)
        f(); f()
        path = api.last_path()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            f()
        e1.record(); torch.cuda.synchronize()
        res.append((e0.elapsed_time(e1) / iters, path))
    fl = 2.5 * 4.0 * b * h * s * s * d * (0.5 if causal else 1.0)
    print(f"bwd b{b} s{s} h{h}/{hk} d{d} causal={causal} p={p} mask={mask}: arrival order {res[0][0]:.3f} ms {fl / res[0][0] / 1e9:.1f} TFLOP/s (path {res[0][1]}), "
          f"deterministic {res[1][0]:.3f} ms {fl / res[1][0] / 1e9:.1f} TFLOP/s (path {res[1][1]})", flush=True)

run(16, 4096, 32, 32, 128, True, 0.0)
run(16, 4096, 32, 32, 128, False, 0.0)
run(16, 4096, 32, 32, 128, True, 0.1)
run(16, 4096, 32, 8, 128, True, 0.0)
run(16, 4096, 32, 32, 64, True, 0.0)
run(8, 2048, 16, 16, 128, True, 0.0)
run(8, 4096, 32, 32, 128, False, 0.0, mask=True)
