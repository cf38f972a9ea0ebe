"""Device-timed backward (and forward) of head sizes above 128 through the C ABI."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

def run(b, s, h, hk, d, causal, p, path=api.PATH_AUTO, iters=10, det=False, mask=False):
    q, do = (torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16) for _ in range(2))
    k, v = (torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16) for _ in range(2))
    m = torch.randn(1, 1, s, s, device="cuda", dtype=torch.bfloat16) if mask else None
    kw = dict(p_dropout=p, is_causal=causal, seed=5, offset=1, attn_mask=m)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, **kw)
    fpath = api.last_path()
    sq_r, d_r = api.round_up(s, 128), api.round_up(d, 32)
    scratch = (torch.empty((b, h, sq_r), dtype=torch.float32, device="cuda"), torch.empty((b, sq_r, h, d_r), dtype=torch.float32, device="cuda"))
    grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    api.set_path(path)
    try:
        f = lambda: api.flash_attn_bwd(do, q, k, v, out, lse, scratch=scratch, grads=grads, deterministic=det, **kw)
        f(); f()
        bpath = api.last_path()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            f()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
    finally:
        api.set_path(api.PATH_AUTO)
    fl = 2.5 * 4.0 * b * h * s * s * d * (0.5 if causal else 1.0)
    print(f"bwd b{b} s{s} h{h}/{hk} d{d} causal={causal} p={p}{' deterministic' if det else ''}{' attn_mask' if mask else ''} paths={fpath}/{bpath}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s", flush=True)

run(4, 2048, 16, 16, 256, True, 0.0)
run(4, 2048, 16, 16, 256, True, 0.0, path=api.PATH_GENERIC, iters=3)
run(4, 2048, 16, 16, 256, False, 0.0)
run(4, 2048, 16, 16, 256, True, 0.1)
run(4, 2048, 16, 16, 192, True, 0.0)
run(8, 4096, 16, 4, 256, True, 0.0)
run(4, 2048, 16, 16, 160, True, 0.0)

if "det" in sys.argv:
    run(4, 2048, 16, 16, 256, True, 0.0, det=True)
    run(4, 2048, 16, 16, 256, False, 0.0, det=True)
    run(8, 4096, 16, 4, 256, True, 0.0, det=True)
    run(4, 2048, 16, 16, 256, False, 0.0, mask=True)
    run(4, 2048, 16, 16, 192, True, 0.1, det=True)
