"""A/B of the host-buffer pipeline (flash_attention_hip_b200/host_pipeline.py) on bench.py's workload: one event per chunk
and direction against the two-event form (forward starts when q, k, v are in; `out` leaves while the backward runs)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200.host_pipeline import HostPipelinedAttention

B, S, H, D = 16, 4096, 32, 128
pin = lambda: torch.empty((B, S, H, D), dtype=torch.bfloat16).pin_memory()
hq, hk, hv, hdo = (pin().normal_() for _ in range(4))
ho, hdq, hdk, hdv = (pin() for _ in range(4))
kw = dict(causal=True, p_dropout=0.1, seed=1, offset=2)
pipes = {fine: HostPipelinedAttention(B, S, H, H, D, chunks=16, fine_events=fine) for fine in (False, True)}
for p in pipes.values():
    p.step(hq, hk, hv, hdo, ho, hdq, hdk, hdv, **kw)
torch.cuda.synchronize()
for rnd in range(3):
    for fine, p in pipes.items():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            p.step(hq, hk, hv, hdo, ho, hdq, hdk, hdv, **kw)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"round {rnd} fine_events={fine}: {ms:.2f} ms per step, {2 * 2**30 / ms / 1e6:.1f} GB/s per direction", flush=True)
