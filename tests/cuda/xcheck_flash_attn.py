"""One-off cross-check against an INDEPENDENT implementation: the upstream flash_attn 2.8 wheel on the GPU image
("secondary cross-check, optional and never the oracle").  Dropout off (its Philox mapping
is its own).  Prints, per shape, the max-abs difference of O / LSE / dQ / dK / dV between this library and the wheel,
relative to the tensor's max-abs, beside each implementation's distance to fp32 math on sampled units.
  python tests/cuda/xcheck_flash_attn.py > gpurun_out/xcheck_flash_attn.log"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from flash_attention_hip_b200 import api

try:
    from flash_attn import flash_attn_func, flash_attn_varlen_func
    import flash_attn
    print("flash_attn wheel", flash_attn.__version__)
except Exception as exc:  # noqa: BLE001
    print("flash_attn wheel unavailable:", exc)
    sys.exit(0)


def rel(a, b):
    return float((a.float() - b.float()).abs().max() / b.float().abs().max())


def unit_ref(q, k, v, do, scale, causal):
    q, k, v, do = (t.float() for t in (q, k, v, do))
    s = (q @ k.t()) * scale
    if causal:
        sq, sk = s.shape
        s = s.masked_fill(torch.ones(sq, sk, dtype=torch.bool, device=s.device).triu(1 + sk - sq), float("-inf"))
    lse = torch.logsumexp(s, -1)
    p = torch.exp(s - lse[:, None])
    out = p @ v
    dv = p.t() @ do
    dp = do @ v.t()
    ds = p * (dp - (do * out).sum(-1, keepdim=True))
    return out, lse, scale * ds @ k, scale * ds.t() @ q, dv


def dense(b, sq, sk, h, hk, d, causal, dtype=torch.bfloat16):
    g = torch.Generator(device="cuda").manual_seed(b * 1000 + sq + d)
    mk = lambda *s: torch.randn(s, generator=g, device="cuda", dtype=torch.float32).to(dtype)
    q, k, v, do = mk(b, sq, h, d), mk(b, sk, hk, d), mk(b, sk, hk, d), mk(b, sq, h, d)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=causal)
    dq, dk, dv = api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=causal)
    qq, kk, vv = (t.clone().requires_grad_(True) for t in (q, k, v))
    o2, lse2, _ = flash_attn_func(qq, kk, vv, causal=causal, return_attn_probs=True)
    o2.backward(do)
    line = (f"dense b{b} sq{sq} sk{sk} h{h}/{hk} d{d} causal={int(causal)} {str(dtype)[6:]}: ours-vs-wheel  O {rel(out, o2):.2e}  "
            f"LSE(abs) {float((lse - lse2[:, :, :sq]).abs().max()):.2e}  dQ {rel(dq, qq.grad):.2e}  dK {rel(dk, kk.grad):.2e}  dV {rel(dv, vv.grad):.2e}")
    if h == hk:
        r = unit_ref(q[0, :, 0], k[0, :, 0], v[0, :, 0], do[0, :, 0], d ** -0.5, causal)
        ours = (out[0, :, 0], dq[0, :, 0], dk[0, :, 0], dv[0, :, 0])
        theirs = (o2[0, :, 0], qq.grad[0, :, 0], kk.grad[0, :, 0], vv.grad[0, :, 0])
        refs = (r[0], r[2], r[3], r[4])
        line += "   | vs fp32 unit (b0,h0): ours " + " ".join(f"{rel(a, x):.2e}" for a, x in zip(ours, refs)) + \
                "  wheel " + " ".join(f"{rel(a, x):.2e}" for a, x in zip(theirs, refs))
    print(line, flush=True)


def packed(lens, h, hk, d, causal):
    import itertools
    cu = torch.tensor([0] + list(itertools.accumulate(lens)), dtype=torch.int32, device="cuda")
    total, mx = sum(lens), max(lens)
    g = torch.Generator(device="cuda").manual_seed(total + d)
    mk = lambda *s: torch.randn(s, generator=g, device="cuda", dtype=torch.float32).bfloat16()
    q, k, v, do = mk(total, h, d), mk(total, hk, d), mk(total, hk, d), mk(total, h, d)
    out, lse, _, _ = api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, is_causal=causal)
    dq, dk, dv = api.flash_attn_varlen_bwd(do, q, k, v, out, lse, cu, cu, mx, mx, is_causal=causal)
    qq, kk, vv = (t.clone().requires_grad_(True) for t in (q, k, v))
    o2 = flash_attn_varlen_func(qq, kk, vv, cu, cu, mx, mx, causal=causal)
    o2.backward(do)
    print(f"packed C4 mix {len(lens)} seqs {total} tokens h{h}/{hk} d{d} causal={int(causal)}: ours-vs-wheel  O {rel(out, o2):.2e}  "
          f"dQ {rel(dq, qq.grad):.2e}  dK {rel(dk, kk.grad):.2e}  dV {rel(dv, vv.grad):.2e}", flush=True)


if __name__ == "__main__":
    dense(1, 128, 128, 4, 4, 64, False)                 # BASELINE.json config 1
    dense(8, 2048, 2048, 16, 16, 128, True)             # config 2
    dense(4, 4096, 4096, 32, 32, 128, True)             # config 3's shape (dropout off), 4 of 16 batches
    dense(2, 1000, 1500, 8, 2, 128, True)
    dense(2, 777, 777, 8, 8, 64, False, torch.float16)
    dense(2, 512, 512, 4, 4, 96, True)
    dense(2, 512, 512, 4, 2, 256, True)
    C4 = [128, 8192, 512, 4096, 1000, 2048, 256, 6000]
    for d_ in (128, 64):
        for c_ in (True, False):
            packed(C4, 32, 8, d_, c_)
