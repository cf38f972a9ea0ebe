"""One wide-head (D=256) backward and one deterministic (two-launch) D=128 backward for ncu captures:
launch order = tc_bwd_wide_kernel (warm-up), tc_bwd_kernel<mode 1>, <mode 2> (warm-up), then the same three again."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

def case(b, s, h, d, det):
    q, k, v, do = (torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16) for _ in range(4))
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=True)
    return lambda: api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=True, deterministic=det)

wide = case(4, 2048, 16, 256, False)
det = case(16, 4096, 32, 128, True)
for _ in range(2):
    wide(); det()
torch.cuda.synchronize()
print("ok", api.last_path())
