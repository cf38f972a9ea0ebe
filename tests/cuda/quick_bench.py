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

WHICH = sys.argv[1] if len(sys.argv) > 1 else "all"

if __name__ == "__main__" and WHICH in ("all", "fwd"):
    time_fwd(8, 2048, 16, 128, True)
    time_fwd(16, 4096, 32, 128, True)
    time_fwd(16, 4096, 32, 128, True, p=0.1)
    time_fwd(16, 4096, 32, 128, False)
    time_fwd(4, 8192, 16, 128, False)
    time_fwd(16, 4096, 32, 64, True)

if __name__ == "__main__" and WHICH == "fdrop":
    for _ in range(2):
        time_fwd(16, 4096, 32, 128, True, p=0.1)
        time_fwd(16, 4096, 32, 128, False, p=0.1)
    time_fwd(8, 2048, 16, 128, True, p=0.1)
    time_fwd(16, 4096, 32, 64, True, p=0.1)

if __name__ == "__main__" and WHICH == "fscan":   # fixed cost per CTA vs cost per KV tile, dropout forward
    for sl, bb in ((512, 1024), (1024, 256), (2048, 64), (4096, 16)):
        time_fwd(bb, sl, 32, 128, False, p=0.1)
    for sl, bb in ((512, 1024), (1024, 256), (2048, 64), (4096, 16)):
        time_fwd(bb, sl, 32, 128, False, p=0.0)

if __name__ == "__main__" and WHICH == "fscanc":
    for sl, bb in ((512, 1024), (1024, 256), (2048, 64), (4096, 16), (8192, 4), (16384, 1)):
        time_fwd(bb, sl, 32, 128, True, p=0.1)


def time_bwd(b, s, h, d, causal, p=0.0, iters=10, dtype=torch.bfloat16, hk=None):
    hk = hk or h
    q = torch.randn(b, s, h, d, device="cuda", dtype=dtype)
    k = torch.randn(b, s, hk, d, device="cuda", dtype=dtype); v = torch.randn_like(k)
    do = torch.randn_like(q)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=causal, p_dropout=p)
    sq_r, d_r = api.round_up(s, 128), api.round_up(d, 32)
    scratch = (torch.empty((b, h, sq_r), dtype=torch.float32, device="cuda"),
               torch.empty((b, sq_r, h, d_r), dtype=torch.float32, device="cuda"))
    grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    for _ in range(2):
        api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=causal, p_dropout=p, scratch=scratch, grads=grads)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=causal, p_dropout=p, scratch=scratch, grads=grads)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = 2.5 * 4.0 * b * h * s * s * d * (0.5 if causal else 1.0)
    print(f"bwd b{b} s{s} h{h}/{hk} d{d} causal={causal} p={p} path={api.last_path()}: {ms:.3f} ms  {fl / ms / 1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__" and WHICH in ("all", "bwd"):
    time_bwd(8, 2048, 16, 128, True)
    time_bwd(16, 4096, 32, 128, True)
    time_bwd(16, 4096, 32, 128, True, p=0.1)
    time_bwd(16, 4096, 32, 128, False)
    time_bwd(16, 4096, 32, 64, True)
    time_bwd(4, 8192, 32, 128, True, hk=8)

if __name__ == "__main__" and WHICH == "bdrop":
    for _ in range(2):
        time_bwd(16, 4096, 32, 128, True, p=0.1)
        time_bwd(16, 4096, 32, 128, False, p=0.1)
    time_bwd(8, 2048, 16, 128, True, p=0.1)
    time_bwd(16, 4096, 32, 64, True, p=0.1)
    time_bwd(4, 8192, 32, 128, True, p=0.1, hk=8)

if __name__ == "__main__" and WHICH == "bigd":
    for d in (160, 192, 256):
        time_fwd(4, 2048, 16, d, True)
        time_bwd(4, 2048, 16, d, True)


def time_membound(b, sq, sk, h, d, causal=False, iters=50, hk=None):
    """Short-sequence / decode-like shapes are HBM- or latency-bound: report achieved GB/s on the
    algorithmic bytes (read Q, K, V; write O, LSE)."""
    hk = hk or h
    q = torch.randn(b, sq, h, d, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(b, sk, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k)
    out = torch.empty_like(q)
    for _ in range(5):
        api.flash_attn_fwd(q, k, v, is_causal=causal, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        api.flash_attn_fwd(q, k, v, is_causal=causal, out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    nbytes = 2 * (2 * q.numel() + 2 * k.numel()) + 4 * b * h * sq
    fl = 4.0 * b * h * sq * sk * d * (0.5 if causal and sq == sk else 1.0)
    print(f"fwd-membound b{b} sq{sq} sk{sk} h{h}/{hk} d{d} causal={causal} path={api.last_path()}: {ms * 1e3:.1f} us  "
          f"{nbytes / ms / 1e6:.1f} GB/s  ({fl / ms / 1e9:.1f} TFLOP/s, {nbytes / 1e6:.2f} MB)", flush=True)


if __name__ == "__main__" and WHICH in ("all", "mem"):
    time_membound(1, 128, 128, 4, 64)                 # BASELINE.json config 1
    time_membound(256, 128, 128, 16, 128)             # many short sequences
    time_membound(64, 256, 256, 32, 128, causal=True)
    time_membound(64, 1, 4096, 32, 128, hk=8)         # decode-like: one query row per sequence, GQA
    time_membound(16, 1, 16384, 32, 128, hk=8)


def time_mask(b, s, h, d, iters=10):
    """attn_mask path (additive bias [1, 1, s, s] broadcast over batch and heads) on the tcgen05 kernels."""
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
    k = torch.randn_like(q); v = torch.randn_like(q); do = torch.randn_like(q)
    mask = torch.randn(1, 1, s, s, device="cuda", dtype=torch.bfloat16)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, attn_mask=mask)
    pf = api.last_path()
    api.flash_attn_bwd(do, q, k, v, out, lse, attn_mask=mask)
    pb_ = api.last_path()
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    for _ in range(iters):
        api.flash_attn_fwd(q, k, v, attn_mask=mask, out=out)
    e[1].record()
    for _ in range(iters):
        api.flash_attn_bwd(do, q, k, v, out, lse, attn_mask=mask)
    e[2].record(); torch.cuda.synchronize()
    fl = 4.0 * b * h * s * s * d
    tf, tb = e[0].elapsed_time(e[1]) / iters, e[1].elapsed_time(e[2]) / iters
    print(f"attn_mask b{b} s{s} h{h} d{d} paths={pf}/{pb_}: fwd {tf:.3f} ms {fl / tf / 1e9:.1f} TFLOP/s, "
          f"bwd {tb:.3f} ms {2.5 * fl / tb / 1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__" and WHICH in ("all", "mask"):
    time_mask(8, 4096, 32, 128)


if __name__ == "__main__" and WHICH in ("all", "det"):
    import functools
    _orig = api.flash_attn_bwd
    api.flash_attn_bwd = functools.partial(_orig, deterministic=True)
    print("deterministic dQ (num_splits == 1):")
    time_bwd(16, 4096, 32, 128, True)
    time_bwd(16, 4096, 32, 128, False)
    api.flash_attn_bwd = _orig


def time_varlen(lens, h, hk, d, causal, iters=10):
    """BASELINE.json config 4: packed mixed-length batch through flash_attn_varlen_fwd / _bwd."""
    import itertools
    cu = torch.tensor([0] + list(itertools.accumulate(lens)), dtype=torch.int32, device="cuda")
    total, mx = sum(lens), max(lens)
    q = torch.randn(total, h, d, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(total, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k)
    do = torch.randn_like(q)
    out, lse, _, _ = api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, is_causal=causal)
    pf = api.last_path()
    api.flash_attn_varlen_bwd(do, q, k, v, out, lse, cu, cu, mx, mx, is_causal=causal)
    pb_ = api.last_path()
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    for _ in range(iters):
        api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, is_causal=causal)
    e[1].record()
    for _ in range(iters):
        api.flash_attn_varlen_bwd(do, q, k, v, out, lse, cu, cu, mx, mx, is_causal=causal)
    e[2].record(); torch.cuda.synchronize()
    fl = sum(4.0 * h * s * s * d * (0.5 if causal else 1.0) for s in lens)
    tf, tb = e[0].elapsed_time(e[1]) / iters, e[1].elapsed_time(e[2]) / iters
    print(f"varlen C4 mix {len(lens)} seqs {total} tokens h{h}/{hk} d{d} causal={causal} paths={pf}/{pb_}: "
          f"fwd {tf:.3f} ms {fl / tf / 1e9:.1f} TFLOP/s, bwd {tb:.3f} ms {2.5 * fl / tb / 1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__" and WHICH in ("all", "varlen"):
    C4 = [128, 8192, 512, 4096, 1000, 2048, 256, 6000]
    for d_ in (128, 64):
        for c_ in (True, False):
            time_varlen(C4, 32, 8, d_, c_)
    time_varlen([128, 256] * 64, 32, 8, 128, True)      # the short end alone
