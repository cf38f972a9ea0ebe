"""Four small calls for compute-sanitizer: dense dropout fwd+bwd, packed, additive mask,
deterministic dQ (num_splits == 1) -- every tcgen05 kernel family, several work items per persistent CTA kept small.
  compute-sanitizer --tool memcheck|racecheck|synccheck python tests/cuda/sanitize_cases.py
One tool per gpurun call (B200_PROFILING.md).  Prints 'sanitize_cases ok' when every call returned."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from flash_attention_hip_b200 import api

torch.manual_seed(0)
mk = lambda *s: torch.randn(s, device="cuda", dtype=torch.float32).bfloat16()


def dense(b, s, h, hk, d, **kw):
    q, k, v, do = mk(b, s, h, d), mk(b, s, hk, d), mk(b, s, hk, d), mk(b, s, h, d)
    det = kw.pop("deterministic", False)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, **kw)
    assert api.last_path() == api.PATH_TCGEN05
    g = api.flash_attn_bwd(do, q, k, v, out, lse, deterministic=det, **kw)
    assert api.last_path() == api.PATH_TCGEN05
    torch.cuda.synchronize()
    assert all(bool(torch.isfinite(t.float()).all()) for t in (out,) + tuple(g))


def packed(lens, h, hk, d, **kw):
    import itertools
    cu = torch.tensor([0] + list(itertools.accumulate(lens)), dtype=torch.int32, device="cuda")
    t, mx = sum(lens), max(lens)
    q, k, v, do = mk(t, h, d), mk(t, hk, d), mk(t, hk, d), mk(t, h, d)
    out, lse, _, _ = api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, **kw)
    assert api.last_path() == api.PATH_TCGEN05
    api.flash_attn_varlen_bwd(do, q, k, v, out, lse, cu, cu, mx, mx, **kw)
    assert api.last_path() == api.PATH_TCGEN05
    torch.cuda.synchronize()


which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "dropout"):
    dense(2, 384, 4, 2, 128, is_causal=True, p_dropout=0.1, seed=1234, offset=7)
    dense(1, 300, 2, 2, 64, p_dropout=0.5, seed=3, offset=2**33)
if which in ("all", "packed"):
    packed([200, 77, 300], 4, 2, 128, is_causal=True, p_dropout=0.1, seed=5)
if which in ("all", "mask"):
    dense(2, 256, 2, 2, 128, attn_mask=mk(1, 2, 256, 256))
if which in ("all", "det"):
    dense(2, 384, 4, 2, 128, is_causal=True, deterministic=True)
if which in ("all", "persistent"):          # more items than SMs: the item queue, the self-resetting counters
    dense(20, 512, 8, 8, 64, is_causal=True, p_dropout=0.1, seed=9)
print("sanitize_cases ok", which)
