"""Multi-GPU check + timing of the ring context-parallel driver (BASELINE.json config 5).
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/cuda/ring_check.py [S_total] [check]
Parity (when `check` is 1 and S_total <= 32768): rank 0 also computes full-sequence attention with the
plain entry points on its own GPU and every rank compares its rows (2e-2 relative, LSE 1e-3).
Timing: device events, max over ranks; flops = causal full-sequence fwd / 2.5x bwd."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import torch.distributed as dist
from flash_attention_hip_b200 import api, ring

S = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
CHECK = (int(sys.argv[2]) if len(sys.argv) > 2 else 1) and S <= 32768
B, H, HK, D = 1, 32, 8, 128
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
g = torch.Generator(device="cuda").manual_seed(4)            # same stream on every rank
full = [torch.randn(B, S, hh, D, generator=g, device="cuda", dtype=torch.float32).bfloat16() for hh in (H, HK, HK, H)]
ql, kl, vl, dol = (ring.zigzag_shard(x, rank, world) for x in full)
if not CHECK:
    del full
    torch.cuda.empty_cache()

def step():
    out, lse = ring.ring_attn_fwd(ql, kl, vl, causal=True)
    grads = ring.ring_attn_bwd(dol, ql, kl, vl, out, lse, causal=True)
    return out, lse, grads

out, lse, (dq, dk, dv) = step()
torch.cuda.synchronize()
if CHECK:
    q, k, v, do = full
    ro, rl, _, _ = api.flash_attn_fwd(q, k, v, is_causal=True)
    rdq, rdk, rdv = api.flash_attn_bwd(do, q, k, v, ro, rl, is_causal=True)
    rel = lambda a, r: float((a.float() - r.float()).abs().max() / r.float().abs().max())
    z = lambda x, dim=1: ring.zigzag_shard(x, rank, world, dim)
    errs = dict(out=rel(out, z(ro)), dq=rel(dq, z(rdq)), dk=rel(dk, z(rdk)), dv=rel(dv, z(rdv)),
                lse=float((lse - z(rl, 2)).abs().max()))
    ok = all(errs[n] < 2e-2 for n in ("out", "dq", "dk", "dv")) and errs["lse"] < 1e-3
    print(f"rank {rank}: parity {'OK' if ok else 'FAIL'} {errs}", flush=True)
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    if int(flag.item()):
        sys.exit(1)

if len(sys.argv) > 3 and sys.argv[3] == "notime":
    ring.close_comms()
    dist.destroy_process_group()
    sys.exit(0)


def timed(fn, iters=3):
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())

t_fwd = timed(lambda: ring.ring_attn_fwd(ql, kl, vl, causal=True))
t_all = timed(step)
fwd_flops = 4.0 * B * H * S * S * D * 0.5
if rank == 0:
    print(json.dumps({"workload": f"C5 ring CP causal B={B} H={H}/{HK} S={S} D={D} bf16, zig-zag", "n_gpus": world,
                      "fwd_ms": t_fwd, "fwd_tflops_total": fwd_flops / t_fwd / 1e9,
                      "fwd_bwd_ms": t_all, "fwd_bwd_tflops_total": 3.5 * fwd_flops / t_all / 1e9,
                      "per_gpu_fwd_tflops": fwd_flops / t_fwd / 1e9 / world, "engine": ring.ENGINE}), flush=True)
ring.close_comms()
dist.destroy_process_group()
