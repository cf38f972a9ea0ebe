"""Per-step device times of the bench workload right after a cold start (clock ramp / power cap study)."""
import os, sys, subprocess, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api
b, h, s, d = 16, 32, 4096, 128
q, k, v, do = (torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16) for _ in range(4))
out = torch.empty_like(q); grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
scratch = (torch.empty((b, h, s), dtype=torch.float32, device="cuda"), torch.empty((b, s, h, d), dtype=torch.float32, device="cuda"))
kw = dict(p_dropout=0.1, is_causal=True, seed=1234, offset=0)
def step():
    _, lse, _, _ = api.flash_attn_fwd(q, k, v, out=out, **kw)
    api.flash_attn_bwd(do, q, k, v, out, lse, scratch=scratch, grads=grads, **kw)
smi = None
if len(sys.argv) > 1 and sys.argv[1] == "smi":
    smi = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "50"],
                           stdout=open("gpurun_out/series_clocks.csv", "w"))
torch.cuda.synchronize(); time.sleep(1.0)
n = 120
ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
ev[0].record()
for i in range(n):
    step(); ev[i + 1].record()
torch.cuda.synchronize()
ts = [ev[i].elapsed_time(ev[i + 1]) for i in range(n)]
print("per-step ms:", " ".join(f"{t:.1f}" for t in ts))
if smi: smi.terminate()
