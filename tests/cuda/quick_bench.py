"""Quick device-side timing of the forward (not the contract bench; see bench.py)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

def time_fwd(b, s, h, d, causal, p=0.0, iters=20, dtype=torch.bfloat16):
    q = torch.randn(b, s, h, d, device="cuda", dtype=dtype)
    k = torch.randn_like(q); v = torch.randn_like(q)
    out = torch.empty_like(q)
    for _ in range(3):
        api.flash_attn_fwd(q, k, v, is_causal=causal, p_dropout=p, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        api.flash_attn_fwd(q, k, v, is_causal=causal, p_dropout=p, out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = 4.0 * b * h * s * s * d * (0.5 if causal else 1.0)
    print(f"fwd b{b} s{s} h{h} d{d} causal={causal} p={p} path={api.last_path()}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s", flush=True)

if __name__ == "__main__":
    time_fwd(8, 2048, 16, 128, True)
    time_fwd(16, 4096, 32, 128, True)
    time_fwd(16, 4096, 32, 128, True, p=0.1)
    time_fwd(16, 4096, 32, 128, False)
    time_fwd(4, 8192, 16, 128, False)
    time_fwd(16, 4096, 32, 64, True)
