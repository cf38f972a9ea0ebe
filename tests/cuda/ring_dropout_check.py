"""Multi-GPU check of dropout in ring mode (flash_attn_b200_ring_fwd_dropout / _bwd_dropout, csrc/fa_ring.cu).
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/cuda/ring_dropout_check.py
Every rank compares its rows of the C engine's result, causal (zig-zag) and not, with
  * tests/ring_ref.py: full-sequence float64 attention under the keep mask assembled from the per-step masks of the
    written specification (2e-2 relative on out / dq / dk / dv, 1e-3 on LSE) -- computed on the host, no ring involved;
  * the Python schedule of ring.py over torch.distributed P2P with the ordinary entry points per step (same kernels, other
    host code: 5e-3)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist
from flash_attention_hip_b200 import ring
from flash_attention_hip_b200.shard import CudaOps
import ring_ref

B, H, HK, D = 2, 4, 2, 128
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
S = 512 * world
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
g = torch.Generator(device="cuda").manual_seed(12)           # same stream on every rank
full = [torch.randn(B, S, hh, D, generator=g, device="cuda", dtype=torch.float32).bfloat16() for hh in (H, HK, HK, H)]
kw = dict(p_dropout=0.2, seed=2**41 + 7, offset=2**34 + 3)
scale = D ** -0.5
rel = lambda a, r: float((a.float() - r.float()).abs().max() / r.float().abs().max())
bad = 0
for causal in (True, False):
    if causal:
        loc = lambda x, dim=1: ring.zigzag_shard(x, rank, world, dim)
    else:
        loc = lambda x, dim=1: x.chunk(world, dim=dim)[rank].contiguous()
    ql, kl, vl, dol = (loc(x) for x in full)
    out, lse = ring.ring_attn_fwd(ql, kl, vl, causal=causal, **kw)
    dq, dk, dv = ring.ring_attn_bwd(dol, ql, kl, vl, out, lse, causal=causal, **kw)
    torch.cuda.synchronize()
    ref = ring_ref.full_reference(*(x.float().cpu().numpy() for x in full), scale, causal, world, **kw)
    ref = {n: torch.from_numpy(a).float().cuda() for n, a in ref.items()}
    errs = dict(out=rel(out, loc(ref["out"])), dq=rel(dq, loc(ref["dq"])), dk=rel(dk, loc(ref["dk"])), dv=rel(dv, loc(ref["dv"])),
                lse=float((lse - loc(ref["lse"], 2)).abs().max()))
    ok = all(errs[n] < 2e-2 for n in ("out", "dq", "dk", "dv")) and errs["lse"] < 1e-3
    ops = CudaOps()
    o2, l2 = ring.ring_attn_fwd(ql, kl, vl, causal=causal, ops=ops, **kw)
    g2 = ring.ring_attn_bwd(dol, ql, kl, vl, o2, l2, causal=causal, ops=ops, **kw)
    torch.cuda.synchronize()
    e2 = dict(out=rel(out, o2), dq=rel(dq, g2[0]), dk=rel(dk, g2[1]), dv=rel(dv, g2[2]), lse=float((lse - l2).abs().max()))
    ok2 = all(e2[n] < 5e-3 for n in ("out", "dq", "dk", "dv")) and e2["lse"] < 1e-4
    # p_dropout = 0 through the dropout entry points is the plain ring call
    print(f"rank {rank} causal={causal}: dropout parity {'OK' if ok and ok2 else 'FAIL'} vs written masks {errs} vs python schedule {e2}", flush=True)
    bad += 0 if (ok and ok2) else 1
flag = torch.tensor([bad], device="cuda")
dist.all_reduce(flag)
ring.close_comms()
dist.destroy_process_group()
sys.exit(1 if int(flag.item()) else 0)
