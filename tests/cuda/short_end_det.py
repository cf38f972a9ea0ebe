"""Short packed sequences: the fused backward (with its dq_accum zero + convert passes) against the two-launch pair."""
import itertools, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

def run(lens, h, hk, d, causal, iters=20):
    cu = torch.tensor([0] + list(itertools.accumulate(lens)), dtype=torch.int32, device="cuda")
    total, mx = sum(lens), max(lens)
    q = torch.randn(total, h, d, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(total, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k); do = torch.randn_like(q)
    out, lse, _, _ = api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, is_causal=causal)
    res = []
    for det in (False, True):
        f = lambda: api.flash_attn_varlen_bwd(do, q, k, v, out, lse, cu, cu, mx, mx, is_causal=causal, deterministic=det)
        f(); f(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): f()
        e1.record(); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / iters)
    print(f"packed {len(lens)} seqs max {mx} total {total} h{h}/{hk} d{d} causal={causal}: fused {res[0]:.3f} ms, two-launch {res[1]:.3f} ms", flush=True)

def dense(b, s, h, hk, d, causal, iters=20):
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k); do = torch.randn_like(q)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=causal)
    res = []
    for det in (False, True):
        f = lambda: api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=causal, deterministic=det)
        f(); f(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): f()
        e1.record(); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) / iters)
    print(f"dense b{b} s{s} h{h}/{hk} d{d} causal={causal}: fused {res[0]:.3f} ms, two-launch {res[1]:.3f} ms", flush=True)

run([128, 256] * 64, 32, 8, 128, True)
run([128] * 128, 32, 8, 128, True)
run([512] * 48, 32, 8, 128, True)
run([1024] * 24, 32, 8, 128, True)
dense(256, 128, 16, 16, 128, False)
dense(64, 256, 32, 32, 128, True)
dense(64, 512, 32, 32, 128, True)
dense(32, 1024, 32, 32, 128, True)
dense(256, 128, 16, 16, 64, False)
