#!/usr/bin/env python
"""bench.py -- the round contract benchmark.

Metric (BASELINE.json): attention fwd+bwd TFLOP/s, bf16, seq 4k, d=128.
Workload per GPU: BASELINE.json configs[2] -- B=16 H=32 S=4096 D=128 causal, dropout 0.1,
one flash_attn_fwd + one flash_attn_bwd per step through the C ABI.  With N GPUs every
rank runs that workload on its own batch shard (global batch 16*N, batch-sharded, no
collective): weak scaling.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun ... bench.py --gpus N ...     (one rank per GPU; rank 0 prints the JSON line)

`--impl reference` times the CPU oracle (the reference ships no implementation of its own,
SURVEY.md section 0) on a bounded sample of the same workload with all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B, H, HK, S, D = 16, 32, 32, 4096, 128
CAUSAL, P_DROP, SEED, OFFSET = True, 0.1, 1234, 0
WORKLOAD = "C3: flash_attn_fwd+flash_attn_bwd bf16 B=16 H=32 S=4096 D=128 causal dropout=0.1 per GPU"


def flops(b, h, s, d, causal):
    fwd = 4.0 * b * h * s * s * d * (0.5 if causal else 1.0)
    return fwd, 2.5 * fwd


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as fh:
            p = json.load(fh)
        return {"burst": p["bf16_tflops"], "sustained": p["bf16_tflops_sustained"], "hbm": p["hbm_gbs"],
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def load_traffic():
    """DRAM bytes per launch of the dominant kernel (dram__bytes_read.sum + dram__bytes_write.sum) from the
    committed ncu --set full capture of this workload (profiles/roofline_traffic.json); None if absent."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(path) as fh:
            return json.load(fh)["tc_bwd_kernel"]["dram_bytes_per_launch"]
    except Exception:
        return None


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons of one GPU with nvidia-smi while the timed region runs."""

    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    # Started BEFORE the warm-up (forking nvidia-smi from a process with a CUDA context stalls the
    # launching thread for tens of ms); only samples read inside [t_begin, t_end] are summarised.

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag, self.proc = index, [], False, None
        self.t_begin, self.t_end = 0.0, float("inf")

    def run(self):
        # one streaming nvidia-smi (a sample every 50 ms) instead of one process per sample
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self.stop_flag:
                    break
                if line.strip():
                    self.samples.append([time.time()] + [x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        self.stop_flag = True
        if self.proc is not None:
            try:
                self.proc.terminate()
            except Exception:
                pass
        self.join(timeout=2)

    def summary(self):
        rows = [r[1:] for r in self.samples if self.t_begin - 0.06 <= r[0] <= self.t_end + 0.06]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(int(r[0]) for r in rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": int(rows[0][1]) if rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(rows)}


def cpu_sample(steps=1, heads=None):
    """Times the CPU oracle (fwd+bwd) on a bounded sample: B=1, H=threads, same S/D/causal/dropout."""
    import numpy as np
    from oracle import oracle as O
    nthreads = O.num_threads()
    h = heads or max(1, min(nthreads, 64))
    pb = O.Problem(batch=1, seqlen_q=S, seqlen_k=S, num_heads=h, num_heads_k=h, head_size=D, is_causal=CAUSAL,
                   p_dropout=P_DROP, seed=SEED, offset=OFFSET)
    rng = np.random.default_rng(2)
    q, k, v, do = (O.to_storage(rng.standard_normal((S, h, D), dtype=np.float32), True) for _ in range(4))
    best = None
    for _ in range(steps):
        t0 = time.perf_counter()
        f = O.forward(pb, q, k, v)
        O.backward(pb, do, q, k, v, f["out"], f["lse"])
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    ff, fb = flops(1, h, S, D, CAUSAL)
    return {"value": (ff + fb) / best / 1e12, "unit": "TFLOP/s", "cores": nthreads, "kind": "port",
            "sample": f"oracle fwd+bwd, B=1 H={h} S={S} D={D} causal dropout={P_DROP}, {best:.2f} s, "
                      f"{nthreads} OpenMP threads; full step = {B * H / h:.0f}x this sample",
            "seconds": best}


def run_reference(args, rank, world):
    if rank != 0:
        return
    # torchrun pins OMP_NUM_THREADS=1 for its workers; the CPU arm is meant to use every host core
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    vals = []
    last = None
    for i in range(args.warmup + args.steps):
        r = cpu_sample(steps=1)
        if i >= args.warmup:
            vals.append(r)
        last = r
    secs = sum(v["seconds"] for v in vals) / len(vals)
    ff, fb = flops(1, int(last["sample"].split("H=")[1].split()[0]), S, D, CAUSAL)
    val = (ff + fb) / secs / 1e12
    line = {"impl": "reference", "metric": "attention fwd+bwd TFLOP/s (bf16 storage, seq 4k, d=128)", "value": val,
            "unit": "TFLOP/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": secs * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": last["sample"]},
            "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": last["cores"], "kind": "port",
                             "sample": last["sample"]},
            "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from flash_attention_hip_b200 import api

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    api.lib()
    peaks = load_peaks()

    g = torch.Generator(device=dev).manual_seed(2 + rank)
    shape = (B, S, H, D)
    q, k, v, dout = (torch.randn(shape, device=dev, dtype=torch.bfloat16, generator=g) for _ in range(4))
    out = torch.empty_like(q)
    grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    sq_r, d_r = api.round_up(S, 128), api.round_up(D, 32)
    scratch = (torch.empty((B, H, sq_r), dtype=torch.float32, device=dev),
               torch.empty((B, sq_r, H, d_r), dtype=torch.float32, device=dev))
    kw = dict(p_dropout=P_DROP, is_causal=CAUSAL, seed=SEED, offset=OFFSET)
    api.set_batch_base(rank * B)      # this rank's batches inside the global batch: shard-invariant dropout
    launches = [0]

    def fwd():
        r = api.flash_attn_fwd(q, k, v, out=out, **kw)
        launches[0] += api.last_launch_count()
        return r[1]

    def bwd(lse):
        api.flash_attn_bwd(dout, q, k, v, out, lse, scratch=scratch, grads=grads, **kw)
        launches[0] += api.last_launch_count()

    def step():
        bwd(fwd())

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)                   # let the fork finish before anything is timed
    for _ in range(max(args.warmup, 5)):
        step()
    fwd_path = None
    lse = fwd(); fwd_path = api.last_path()
    bwd(lse); bwd_path = api.last_path()
    sync_all()

    launches[0] = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    sampler.t_begin = time.time()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    sync_all()
    sampler.t_end = time.time()
    ms = e0.elapsed_time(e1)
    n_launch = launches[0]

    # per-kernel-family timings for the roofline (same stream, CUDA events, after the timed region)
    def time_loop(fn, n, repeats=3):
        # median of `repeats` back-to-back groups: one group now and then catches a host-side stall (the
        # nvidia-smi sampler holds a driver lock the stream-ordered allocator also wants) and read 2.5x high
        got = []
        for _ in range(repeats):
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record()
            for _ in range(n):
                fn()
            b_.record()
            torch.cuda.synchronize()
            got.append(a.elapsed_time(b_) / n)
        return sorted(got)[len(got) // 2]

    n_k = max(3, min(args.steps, 10))
    fwd_ms = time_loop(fwd, n_k)
    bwd_ms = time_loop(lambda: bwd(lse), n_k)
    sampler.stop()
    clocks = sampler.summary()

    # e2e: host buffers in, host buffers out, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        hq, hk, hv, hdo = (torch.empty(shape, dtype=torch.bfloat16).pin_memory() for _ in range(4))
        for h_, d_ in ((hq, q), (hk, k), (hv, v), (hdo, dout)):
            h_.copy_(d_)
        ho, hdq, hdk, hdv = (torch.empty(shape, dtype=torch.bfloat16).pin_memory() for _ in range(4))
        n_e = max(1, min(args.steps, 3))

        def e2e_sequential():      # copy everything in, run, copy everything out: nothing overlaps
            q.copy_(hq, non_blocking=True); k.copy_(hk, non_blocking=True)
            v.copy_(hv, non_blocking=True); dout.copy_(hdo, non_blocking=True)
            step()
            ho.copy_(out, non_blocking=True); hdq.copy_(grads[0], non_blocking=True)
            hdk.copy_(grads[1], non_blocking=True); hdv.copy_(grads[2], non_blocking=True)

        # the public host-buffer call: batch chunks pipelined over three streams (host_pipeline.py)
        from flash_attention_hip_b200.host_pipeline import HostPipelinedAttention
        pipe = HostPipelinedAttention(B, S, H, H, D, chunks=args.e2e_chunks, device=dev)

        def e2e_step():
            pipe.step(hq, hk, hv, hdo, ho, hdq, hdk, hdv, causal=CAUSAL, p_dropout=P_DROP, seed=SEED, offset=OFFSET,
                      batch_base=rank * B)

        def time_e2e(fn):
            fn()
            sync_all()
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n_e):
                fn()
            b_.record()
            sync_all()
            return a.elapsed_time(b_) / n_e

        e2e_seq_ms = time_e2e(e2e_sequential)
        ref_out = ho.clone()
        e2e_ms = time_e2e(e2e_step)
        if not torch.equal(ref_out, ho):
            raise SystemExit("bench.py: pipelined host call and the one-shot call disagree")
        e2e = (e2e_ms, 4 * q.numel() * 2, 4 * q.numel() * 2, e2e_seq_ms)

    t = torch.tensor([ms, fwd_ms, bwd_ms, e2e[0] if e2e else 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, fwd_ms, bwd_ms, e2e_ms = (float(x) for x in t.cpu())
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ff, fb = flops(B, H, S, D, CAUSAL)
    ms_step = ms / args.steps
    value = world * (ff + fb) / (ms_step * 1e-3) / 1e12
    dom_flops, dom_ms, dom_name = (fb, bwd_ms, "flash_attn_bwd (preprocess + main + postprocess)")
    achieved = dom_flops / (dom_ms * 1e-3) / 1e12
    line = {
        "metric": "attention fwd+bwd TFLOP/s (bf16, seq 4k, d=128)", "value": value, "unit": "TFLOP/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": B * world, "sharding": f"batch x{world}, no collective",
                   "l2": "inputs (5 x 512 MiB per GPU) exceed L2; no flush needed",
                   "fwd_path": {1: "generic", 2: "tcgen05"}.get(fwd_path), "bwd_path": {1: "generic", 2: "tcgen05"}.get(bwd_path)},
        "per_gpu_tflops": value / world,
        "frac_of_peak": {"vs_sustained": value / world / peaks["sustained"], "vs_burst": value / world / peaks["burst"],
                         "vs_nominal_2250": value / world / 2250.0, "peaks": peaks["source"]},
        "kernels": {"fwd_ms": fwd_ms, "fwd_tflops": ff / fwd_ms / 1e9, "bwd_ms": bwd_ms, "bwd_tflops": fb / bwd_ms / 1e9},
        "roofline": {"bound": "tensor", "kernel": dom_name, "achieved": achieved, "peak": peaks["sustained"],
                     "unit": "TFLOP/s", "frac": achieved / peaks["sustained"], "traffic": load_traffic(),
                     "peak_kind": "bf16_tflops_sustained, " + peaks["source"]},
        "roofline_fwd": {"bound": "tensor", "achieved": ff / fwd_ms / 1e9, "peak": peaks["sustained"], "unit": "TFLOP/s",
                         "frac": ff / fwd_ms / 1e9 / peaks["sustained"]},
        "clocks": clocks, "gpu_launches": n_launch,
    }
    if e2e:
        line["e2e"] = {"value": world * (ff + fb) / (e2e_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
                       "h2d_bytes_per_step": e2e[1], "d2h_bytes_per_step": e2e[2], "ms_per_step": e2e_ms,
                       "how": f"pinned host buffers, {args.e2e_chunks} batch chunks pipelined on 3 streams (H2D | kernels | D2H)",
                       "sequential_ms_per_step": e2e[3]}
    if world == 1 and not args.no_cpu:
        try:
            line["cpu_baseline"] = {k_: v_ for k_, v_ in cpu_sample().items() if k_ != "seconds"}
        except Exception as exc:  # the oracle is a checker; its absence must not hide the GPU number
            line["cpu_baseline"] = {"value": None, "error": str(exc)}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=16)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
