"""One fwd+bwd of the bench workload (C3 shape, optionally smaller batch) for ncu captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

b = int(sys.argv[1]) if len(sys.argv) > 1 else 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
h, s, d = 32, 4096, 128
q, k, v, do = (torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16) for _ in range(4))
out = torch.empty_like(q)
grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
scratch = (torch.empty((b, h, s), dtype=torch.float32, device="cuda"),
           torch.empty((b, s, h, d), dtype=torch.float32, device="cuda"))
kw = dict(p_dropout=0.1, is_causal=True, seed=1234, offset=0)
for _ in range(reps):
    _, lse, _, _ = api.flash_attn_fwd(q, k, v, out=out, **kw)
    api.flash_attn_bwd(do, q, k, v, out, lse, scratch=scratch, grads=grads, **kw)
torch.cuda.synchronize()
print("ok", api.last_path())
