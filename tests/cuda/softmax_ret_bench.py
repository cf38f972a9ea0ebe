"""Device time of `return_softmax` through the C ABI: the tcgen05 forward + the S_dmask pass of csrc/fa_softmax_ret.cu
against the same call on the all-shapes kernel.  The wrapper zeroes the plane before every call (torch.zeros); that
fill is timed alone and reported beside the calls."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api


def timed(f, iters):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def run(b, s, h, hk, d, causal, p, generic_iters=1):
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
    k, v = (torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16) for _ in range(2))
    kw = dict(p_dropout=p, is_causal=causal, seed=5, offset=1)
    sr = api.round_up(s, 128)
    plane_gb = b * h * sr * sr * 2 / 1e9
    t_plain = timed(lambda: api.flash_attn_fwd(q, k, v, **kw), 5)
    t_fill = timed(lambda: torch.zeros((b, h, sr, sr), dtype=q.dtype, device="cuda"), 5)
    t_ret = timed(lambda: api.flash_attn_fwd(q, k, v, return_softmax=True, **kw), 5)
    path = api.last_path()
    api.set_path(api.PATH_GENERIC)
    try:
        t_gen = timed(lambda: api.flash_attn_fwd(q, k, v, return_softmax=True, **kw), generic_iters)
    finally:
        api.set_path(api.PATH_AUTO)
    pass_ms = t_ret - t_plain - t_fill
    print(f"return_softmax b{b} s{s} h{h}/{hk} d{d} causal={causal} p={p} path={path}: plane {plane_gb:.2f} GB; forward {t_plain:.3f} ms, "
          f"zero fill {t_fill:.3f} ms, forward + S_dmask pass {t_ret - t_fill:.3f} ms (pass {pass_ms:.3f} ms = {plane_gb / pass_ms * 1e3:.0f} GB/s written); "
          f"all-shapes kernel {t_gen - t_fill:.1f} ms", flush=True)


run(2, 4096, 16, 16, 128, True, 0.1)
run(2, 4096, 16, 4, 128, False, 0.0)
run(4, 2048, 16, 16, 64, True, 0.1)
run(2, 2048, 16, 16, 256, True, 0.0)
