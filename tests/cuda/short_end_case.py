"""The short end of config 4 alone (128 packed sequences of 128 / 256 tokens, H32/8 D128 causal): two forward + backward
calls, for a launch-by-launch breakdown under `ncu --metrics gpu__time_duration.sum`."""
import itertools, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api
lens = [128, 256] * 64
cu = torch.tensor([0] + list(itertools.accumulate(lens)), dtype=torch.int32, device="cuda")
total, mx, h, hk, d = sum(lens), max(lens), 32, 8, 128
q = torch.randn(total, h, d, device="cuda", dtype=torch.bfloat16)
k = torch.randn(total, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k); do = torch.randn_like(q)
for _ in range(2):
    out, lse, _, _ = api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, is_causal=True)
    api.flash_attn_varlen_bwd(do, q, k, v, out, lse, cu, cu, mx, mx, is_causal=True)
torch.cuda.synchronize()
print("ok", api.last_path())
