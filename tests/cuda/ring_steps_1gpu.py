"""Device time of the three step shapes of the causal zig-zag ring at N = 8, S = 131072 (s_local 16384, chunk 8192) on ONE GPU,
and of the world-size-1 ring call (one causal step + the merge kernel) against the plain call: where a ring step's time goes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

def t(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

h, hk, d, s, c = 32, 8, 128, 16384, 8192
q = torch.randn(1, s, h, d, device="cuda", dtype=torch.bfloat16)
k = torch.randn(1, s, hk, d, device="cuda", dtype=torch.bfloat16); v = torch.randn_like(k); do = torch.randn_like(q)
shapes = {"own chunks: causal q[0:2c] x kv[0:2c]": (q, k, v, True),
          "from a lower rank: q[0:2c] x kv[0:c]": (q, k[:, :c].contiguous(), v[:, :c].contiguous(), False),
          "from a higher rank: q[c:2c] x kv[0:2c]": (q[:, c:].contiguous(), k, v, False)}
for name, (qq, kk, vv, causal) in shapes.items():
    out, lse, _, _ = api.flash_attn_fwd(qq, kk, vv, is_causal=causal)
    dd = torch.randn_like(qq)
    fl = 4.0 * h * qq.shape[1] * kk.shape[1] * d * (0.5 if causal else 1.0)
    tf = t(lambda: api.flash_attn_fwd(qq, kk, vv, is_causal=causal))
    tb = t(lambda: api.flash_attn_bwd(dd, qq, kk, vv, out, lse, is_causal=causal))
    print(f"{name}: fwd {tf:.3f} ms {fl / tf / 1e9:.0f} TFLOP/s, bwd {tb:.3f} ms {2.5 * fl / tb / 1e9:.0f} TFLOP/s", flush=True)

import torch.distributed as dist
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29577")
dist.init_process_group("gloo", rank=0, world_size=1)
from flash_attention_hip_b200 import ring
out, lse = ring.ring_attn_fwd(q, k, v, causal=True)
tr = t(lambda: ring.ring_attn_fwd(q, k, v, causal=True))
tp = t(lambda: api.flash_attn_fwd(q, k, v, is_causal=True))
tb_r = t(lambda: ring.ring_attn_bwd(do, q, k, v, out, lse, causal=True))
tb_p = t(lambda: api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=True))
print(f"world-size-1 ring forward {tr:.3f} ms against the plain call {tp:.3f} ms (difference = merge kernel + host schedule); backward {tb_r:.3f} against {tb_p:.3f} ms")
ring.close_comms(); dist.destroy_process_group()
