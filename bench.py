#!/usr/bin/env python
"""bench.py -- the round contract benchmark.

Metric (BASELINE.json): attention fwd+bwd TFLOP/s, bf16, seq 4k, d=128.
Workload per GPU: BASELINE.json configs[2] -- B=16 H=32 S=4096 D=128 causal, dropout 0.1,
one flash_attn_fwd + one flash_attn_bwd per step through the C ABI.  With N GPUs every
rank runs that workload on its own batch shard (global batch 16*N, batch-sharded, no
collective): weak scaling -- the headline line.  The same run also reports, as extra keys:
  "sustained"  the same step looped for >= 3 s (clocks sampled): the fraction of the SUSTAINED peak is
               taken from this loop, the fraction of the BURST peak from the K driver steps;
  "strong"     (N > 1) BASELINE.json configs[2] as worded: B=16 GLOBAL, batch-sharded over the N ranks;
  "ring_c5"    (N > 1) configs[4]: S=131072 causal ring context parallel over NVLink, parity-checked at
               S=16384 against a single-GPU full-sequence run inside the same invocation.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun ... bench.py --gpus N ...     (one rank per GPU; rank 0 prints the JSON line)

`--impl reference` times the CPU oracle (the reference ships no implementation of its own,
SURVEY.md section 0) on a bounded sample of the same workload with all host threads.
"""
from __future__ import annotations

import argparse
import gc
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


BWD_LAUNCHES = ("bwd_preprocess_kernel", "tc_bwd_kernel", "dq_convert_kernel")   # (the preprocess kernel also zeroes dq_accum)


def load_traffic():
    """DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of ONE flash_attn_bwd call on this workload =
    the sum over the launches the call makes, from the committed ncu --set full capture
    (profiles/roofline_traffic.json).  `roofline.kernel` is the same call, so the two fields describe the same
    thing.  None if the capture is absent or lacks one of the launches."""
    path = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        with open(path) as fh:
            t = json.load(fh)
        return sum(t[k]["dram_bytes_per_launch"] for k in t.get("flash_attn_bwd_launches", BWD_LAUNCHES))
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
        # NVML in this process when pynvml is there (a sample every 20 ms, each a few small queries): the streaming
        # nvidia-smi below takes a driver lock long enough to hold a kernel launch back for tens of ms now and then --
        # harmless while the launch queue is full, but the first step after a barrier starts on an empty queue.
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            mx = N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM)
            flag = lambda bits, m: "Active" if bits & m else "Not Active"
            while not self.stop_flag:
                sm = N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)
                r = N.nvmlDeviceGetCurrentClocksEventReasons(h)
                self.samples.append([time.time(), str(sm), str(mx), flag(r, N.nvmlClocksEventReasonHwSlowdown),
                                     flag(r, N.nvmlClocksEventReasonHwThermalSlowdown),
                                     flag(r, N.nvmlClocksEventReasonSwThermalSlowdown),
                                     flag(r, N.nvmlClocksEventReasonSwPowerCap)])
                time.sleep(0.02)
            return
        except Exception:
            if self.samples:
                return
        # fallback: one streaming nvidia-smi (a sample every 50 ms) instead of one process per sample
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


METRIC = "attention fwd+bwd TFLOP/s (bf16, seq 4k, d=128)"     # BASELINE.json's metric; BOTH arms print this string


def run_reference(args, rank, world):
    """The CPU arm: the oracle (a straightforward fp32 row-at-a-time restatement, OpenMP over (batch, head) --
    NOT a tuned CPU attention; the reference ships no implementation of its own) on a bounded sample."""
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
    h_sample = int(last["sample"].split("H=")[1].split()[0])
    ff, fb = flops(1, h_sample, S, D, CAUSAL)
    val = (ff + fb) / secs / 1e12
    full_over_sample = B * H / h_sample
    line = {"impl": "reference", "metric": METRIC, "value": val,
            "unit": "TFLOP/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            # a full step (B=16 H=32) at the sampled rate; the sample itself took sample_ms
            "ms_per_step": secs * 1e3 * full_over_sample, "sample_ms": secs * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": last["sample"],
                       "arithmetic": "bf16 storage, fp32 math (scalar row-at-a-time CPU port, not a tuned CPU attention)"},
            "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": last["cores"], "kind": "port",
                             "sample": last["sample"]},
            "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def _sync_all(torch, dist, world):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


def _max_over_ranks(torch, dist, world, dev, vals):
    t = torch.tensor(list(vals), device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.cpu()]


class Workload:
    """C3 tensors of `b` batches on one device + the fwd / bwd / step closures through the C ABI."""

    def __init__(self, torch, api, dev, b, batch_base, seed):
        g = torch.Generator(device=dev).manual_seed(seed)
        shape = (b, S, H, D)
        self.shape, self.api, self.b = shape, api, b
        self.q, self.k, self.v, self.dout = (torch.randn(shape, device=dev, dtype=torch.bfloat16, generator=g) for _ in range(4))
        self.out = torch.empty_like(self.q)
        self.grads = (torch.empty_like(self.q), torch.empty_like(self.k), torch.empty_like(self.v))
        sq_r, d_r = api.round_up(S, 128), api.round_up(D, 32)
        self.scratch = (torch.empty((b, H, sq_r), dtype=torch.float32, device=dev),
                        torch.empty((b, sq_r, H, d_r), dtype=torch.float32, device=dev))
        self.kw = dict(p_dropout=P_DROP, is_causal=CAUSAL, seed=SEED, offset=OFFSET, batch_base=batch_base)
        self.launches = 0

    def fwd(self):
        r = self.api.flash_attn_fwd(self.q, self.k, self.v, out=self.out, **self.kw)
        self.launches += self.api.last_launch_count()
        return r[1]

    def bwd(self, lse):
        self.api.flash_attn_bwd(self.dout, self.q, self.k, self.v, self.out, lse, scratch=self.scratch, grads=self.grads, **self.kw)
        self.launches += self.api.last_launch_count()

    def step(self):
        self.bwd(self.fwd())


def _ring_c5(torch, dist, api, rank, world, dev):
    """BASELINE.json configs[4]: S=131072 D=128 causal (B=1, H=32/8) ring context parallel over NVLink, zig-zag
    sharded.  Parity first, inside this run: S=16384 over the ring against rank-local full-sequence attention
    (north_star tolerances); then device-timed forward and forward+backward at the full length."""
    from flash_attention_hip_b200 import ring
    hq, hkv, d = 32, 8, 128

    def make(s_total, seed):
        g = torch.Generator(device=dev).manual_seed(seed)          # the same stream on every rank
        return [torch.randn(1, s_total, hh, d, generator=g, device=dev, dtype=torch.float32).bfloat16() for hh in (hq, hkv, hkv, hq)]

    # ---- parity at S = 16384
    full = make(16384, 4)
    ql, kl, vl, dol = (ring.zigzag_shard(x, rank, world) for x in full)
    out, lse = ring.ring_attn_fwd(ql, kl, vl, causal=True)
    dq, dk, dv = ring.ring_attn_bwd(dol, ql, kl, vl, out, lse, causal=True)
    q, k, v, do = full
    ro, rl, _, _ = api.flash_attn_fwd(q, k, v, is_causal=True)
    rdq, rdk, rdv = api.flash_attn_bwd(do, q, k, v, ro, rl, is_causal=True)
    rel = lambda a, r: float((a.float() - r.float()).abs().max() / r.float().abs().max())
    z = lambda x, dim=1: ring.zigzag_shard(x, rank, world, dim)
    errs = [rel(out, z(ro)), rel(dq, z(rdq)), rel(dk, z(rdk)), rel(dv, z(rdv)), float((lse - z(rl, 2)).abs().max())]
    errs = _max_over_ranks(torch, dist, world, dev, errs)
    parity_ok = all(e < 2e-2 for e in errs[:4]) and errs[4] < 1e-3
    del full, q, k, v, do, ro, rl, rdq, rdk, rdv, out, lse, dq, dk, dv, ql, kl, vl, dol
    torch.cuda.empty_cache()

    # ---- timing at S = 131072
    s_total = 131072
    ql, kl, vl, dol = (ring.zigzag_shard(x, rank, world) for x in make(s_total, 5))
    torch.cuda.empty_cache()

    def fwd_only():
        return ring.ring_attn_fwd(ql, kl, vl, causal=True)

    def fwd_bwd():
        o, l = ring.ring_attn_fwd(ql, kl, vl, causal=True)
        ring.ring_attn_bwd(dol, ql, kl, vl, o, l, causal=True)

    def timed(fn, iters):
        fn()
        _sync_all(torch, dist, world)
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            fn()
        b_.record()
        _sync_all(torch, dist, world)
        return _max_over_ranks(torch, dist, world, dev, [a.elapsed_time(b_) / iters])[0]

    t_fwd = timed(fwd_only, 3)
    t_all = timed(fwd_bwd, 2)
    ff = 4.0 * hq * float(s_total) * s_total * d * 0.5
    return {"workload": f"C5: ring context parallel, causal B=1 H={hq}/{hkv} S={s_total} D={d} bf16, zig-zag over {world} GPUs",
            "fwd_ms": t_fwd, "bwd_ms": t_all - t_fwd, "fwd_bwd_ms": t_all,
            "per_gpu_tflops": {"fwd": ff / t_fwd / 1e9 / world, "fwd_bwd": 3.5 * ff / t_all / 1e9 / world},
            "parity_ok": parity_ok,
            "parity": {"at": "S=16384 vs single-GPU full-sequence flash_attn_fwd/bwd, max over ranks",
                       "out": errs[0], "dq": errs[1], "dk": errs[2], "dv": errs[3], "lse_abs": errs[4]},
            "engine": ring.ENGINE if hasattr(ring, "ENGINE") else "python driver over torch.distributed P2P"}


def run_ours(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from flash_attention_hip_b200 import api
    from flash_attention_hip_b200.shard import batch_shard

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    api.lib()
    peaks = load_peaks()

    # this rank's batches inside the global batch (shard-invariant dropout through batch_base)
    w = Workload(torch, api, dev, B, rank * B, 2 + rank)
    sync_all = lambda: _sync_all(torch, dist, world)

    sampler = ClockSampler(local_rank)
    sampler.start()
    # nvidia-smi's start-up (driver attach, first query) stalls kernel launches for tens of ms: once it fell inside
    # the 20 timed steps and cost them 70 ms (12.4 against 8.9 ms per step).  Wait for its first samples instead of
    # a fixed 0.3 s, then warm up.
    t_wait = time.time()
    while len(sampler.samples) < 2 and time.time() - t_wait < 20.0:
        time.sleep(0.05)
    # Warm up for at least W steps AND ~1.5 s: a B200 that was idle runs the first ~60 ms at its 1965 MHz boost
    # (7.9 ms per step here), then the 1 kW software power cap pulls the clock to ~1.5 GHz (9.1-9.3 ms per step);
    # 20 timed steps right after a 5-step warm-up straddle that transition and read anything from 8.9 to 9.9 ms
    # from run to run (profiles/r02e_step_series.log).  Timed after the transition, the K steps agree with the
    # >= 3 s `sustained` loop below.
    t_warm, n_warm = time.time(), 0
    while n_warm < max(args.warmup, 5) or time.time() - t_warm < args.warm_seconds:
        w.step()
        n_warm += 1
        if n_warm % 8 == 0:
            torch.cuda.synchronize()
    lse = w.fwd(); fwd_path = api.last_path()
    w.bwd(lse); bwd_path = api.last_path()
    sync_all()

    # ---- the contract's timed region: K steps ------------------------------------------------------------
    def timed_region():
        """Exactly K steps between two synchronised barriers; per-step event marks show where a slow region lost time."""
        w.launches = 0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
        # The first step after a barrier starts on an EMPTY launch queue, so a host-side pause there idles the GPU.
        # The pauses seen (12-70 ms, `host_enqueue_ms`) were the interpreter's cyclic garbage collector running a full
        # collection over everything torch imports: collect now, and keep it out of the K steps.
        gc.collect()
        gc.disable()
        sync_all()
        sampler.t_begin = time.time()
        t_host0 = time.time()
        e0.record()
        for i_ in range(args.steps):
            w.step()
            marks[i_].record()
        e1.record()
        t_host1 = time.time()
        sync_all()
        sampler.t_end = time.time()          # (the collector stays off for the other timed loops; each collects first)
        steps_ms = [(e0 if i_ == 0 else marks[i_ - 1]).elapsed_time(marks[i_]) for i_ in range(args.steps)]
        return e0.elapsed_time(e1), steps_ms, (t_host1 - t_host0) * 1e3, w.launches, sampler.summary()

    ms, step_ms, host_enqueue_ms, n_launch, clocks = timed_region()
    # A region in which one step took over 1.5x the median step was disturbed from outside the kernels (seen twice on
    # this pool right after another process had used the box: 12.4 and 14.1 ms per step against 9.3 in every other
    # loop of the same run).  Like a throttled run it is measured again, ONCE, and both results are reported.
    remeasured = None
    med = sorted(step_ms)[len(step_ms) // 2]
    disturbed = _max_over_ranks(torch, dist, world, dev, [1.0 if max(step_ms) > 1.5 * med else 0.0])[0] > 0
    if disturbed:
        first = {"ms_per_step": ms / args.steps, "timed_steps_ms": [round(x, 3) for x in step_ms],
                 "host_enqueue_ms": round(host_enqueue_ms, 3)}
        ms, step_ms, host_enqueue_ms, n_launch, clocks = timed_region()
        remeasured = {"reason": "one step of the first K-step region took over 1.5x its median step", "first": first}

    # ---- per-entry-point timings for the roofline (same stream, CUDA events, after the timed region)
    def time_loop(fn, n, repeats=3):
        # median of `repeats` back-to-back groups: one group now and then catches a host-side stall
        got = []
        gc.collect()
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
    fwd_ms = time_loop(w.fwd, n_k)
    bwd_ms = time_loop(lambda: w.bwd(lse), n_k)

    # ---- sustained: the same step looped for >= 3 s (the denominator "sustained peak" is a seconds-long loop too)
    sustained = None
    if not args.no_sustained:
        n_s = max(args.steps, int(args.sustain_seconds * 1e3 / max(ms / args.steps, 1e-3)) + 1)
        gc.collect()
        sync_all()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t_s0 = time.time()
        a.record()
        for _ in range(n_s):
            w.step()
        b_.record()
        sync_all()
        t_s1 = time.time()
        s_ms = a.elapsed_time(b_)
        sampler.t_begin, sampler.t_end = t_s0, t_s1
        sustained = (s_ms, n_s, sampler.summary())
    sampler.stop()

    # ---- strong scaling: configs[2] as worded -- B=16 GLOBAL, batch-sharded over the ranks --------------------
    strong = None
    if world > 1 and not args.no_strong:
        first, count = batch_shard(B, world, rank)
        ws = Workload(torch, api, dev, count, first, 2)      # (rank-local random data; only the shapes matter)
        for _ in range(5):
            ws.step()
        gc.collect()
        sync_all()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            ws.step()
        b_.record()
        sync_all()
        strong = a.elapsed_time(b_)
        del ws
        torch.cuda.empty_cache()

    # ---- e2e: host buffers in, host buffers out, copies inside the timed region ------------------------------
    e2e = None
    if not args.no_e2e:
        shape = w.shape
        hq, hk, hv, hdo = (torch.empty(shape, dtype=torch.bfloat16).pin_memory() for _ in range(4))
        for h_, d_ in ((hq, w.q), (hk, w.k), (hv, w.v), (hdo, w.dout)):
            h_.copy_(d_)
        ho, hdq, hdk, hdv = (torch.empty(shape, dtype=torch.bfloat16).pin_memory() for _ in range(4))
        n_e = max(1, min(args.steps, 3))

        def e2e_sequential():      # copy everything in, run, copy everything out: nothing overlaps
            w.q.copy_(hq, non_blocking=True); w.k.copy_(hk, non_blocking=True)
            w.v.copy_(hv, non_blocking=True); w.dout.copy_(hdo, non_blocking=True)
            w.step()
            ho.copy_(w.out, non_blocking=True); hdq.copy_(w.grads[0], non_blocking=True)
            hdk.copy_(w.grads[1], non_blocking=True); hdv.copy_(w.grads[2], non_blocking=True)

        # the public host-buffer call: batch chunks pipelined over three streams (host_pipeline.py)
        from flash_attention_hip_b200.host_pipeline import HostPipelinedAttention
        pipe = HostPipelinedAttention(B, S, H, H, D, chunks=args.e2e_chunks, device=dev)

        def e2e_step():
            pipe.step(hq, hk, hv, hdo, ho, hdq, hdk, hdv, causal=CAUSAL, p_dropout=P_DROP, seed=SEED, offset=OFFSET,
                      batch_base=rank * B)

        def time_e2e(fn):
            fn()
            gc.collect()
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
        e2e = (e2e_ms, 4 * w.q.numel() * 2, 4 * w.q.numel() * 2, e2e_seq_ms)
        del pipe, hq, hk, hv, hdo, ho, hdq, hdk, hdv, ref_out

    ms, fwd_ms, bwd_ms, e2e_ms, s_ms, strong_ms = _max_over_ranks(
        torch, dist, world, dev, [ms, fwd_ms, bwd_ms, e2e[0] if e2e else 0.0, sustained[0] if sustained else 0.0, strong or 0.0])

    # ---- ring context parallel, config 5 (N > 1) ------------------------------------------------------------
    ring_c5 = None
    if world > 1 and not args.no_ring:
        del w
        torch.cuda.empty_cache()
        try:
            ring_c5 = _ring_c5(torch, dist, api, rank, world, dev)
        except Exception as exc:  # noqa: BLE001 -- reported in the line, never hides the headline
            ring_c5 = {"error": repr(exc)}
        try:
            from flash_attention_hip_b200 import ring as _ring
            _ring.close_comms()
        except Exception:  # noqa: BLE001
            pass

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    ff, fb = flops(B, H, S, D, CAUSAL)
    ms_step = ms / args.steps
    value = world * (ff + fb) / (ms_step * 1e-3) / 1e12
    dom_flops, dom_ms = fb, bwd_ms
    achieved = dom_flops / (dom_ms * 1e-3) / 1e12
    frac = {"vs_burst": value / world / peaks["burst"], "vs_nominal_2250": value / world / 2250.0, "peaks": peaks["source"],
            "note": "vs_burst: the K driver steps (a sub-second loop) against the burst peak; vs_sustained: the "
                    ">= 3 s loop of the same step (key `sustained`) against the sustained peak"}
    line = {
        "metric": METRIC, "value": value, "unit": "TFLOP/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": B * world, "sharding": f"batch x{world}, no collective",
                   "l2": "inputs (5 x 512 MiB per GPU) exceed L2; no flush needed",
                   "fwd_path": {1: "generic", 2: "tcgen05"}.get(fwd_path), "bwd_path": {1: "generic", 2: "tcgen05"}.get(bwd_path)},
        "per_gpu_tflops": value / world,
        "frac_of_peak": frac,
        "kernels": {"fwd_ms": fwd_ms, "fwd_tflops": ff / fwd_ms / 1e9, "bwd_ms": bwd_ms, "bwd_tflops": fb / bwd_ms / 1e9},
        # the dominant entry point (70 % of the step): ONE flash_attn_bwd call = the launches in BWD_LAUNCHES;
        # `achieved` from its CUDA-event time, `traffic` = the DRAM bytes of the same launches summed
        "roofline": {"bound": "tensor", "kernel": "flash_attn_bwd (all launches of one call: " + " + ".join(BWD_LAUNCHES) + ")",
                     "achieved": achieved, "peak": peaks["burst"], "unit": "TFLOP/s", "frac": achieved / peaks["burst"],
                     "frac_vs_sustained_peak": achieved / peaks["sustained"], "traffic": load_traffic(),
                     "algorithmic_bytes": 5 * B * S * H * D * 2 + 3 * B * S * H * D * 2 + B * H * S * 4,
                     "peak_kind": "bf16_tflops (burst: the per-entry-point loops last well under a second), " + peaks["source"]},
        "roofline_fwd": {"bound": "tensor", "kernel": "flash_attn_fwd (tc_fwd_kernel)", "achieved": ff / fwd_ms / 1e9,
                         "peak": peaks["burst"], "unit": "TFLOP/s", "frac": ff / fwd_ms / 1e9 / peaks["burst"],
                         "frac_vs_sustained_peak": ff / fwd_ms / 1e9 / peaks["sustained"]},
        "clocks": clocks, "gpu_launches": n_launch,
        "timed_steps_ms": [round(x, 3) for x in step_ms], "host_enqueue_ms": round(host_enqueue_ms, 3),
    }
    if remeasured:
        line["remeasured"] = remeasured
    if sustained:
        s_step = s_ms / sustained[1]
        s_val = (ff + fb) / (s_step * 1e-3) / 1e12          # per GPU (max over ranks of the loop time)
        line["sustained"] = {"seconds": s_ms * 1e-3, "steps": sustained[1], "ms_per_step": s_step, "per_gpu_tflops": s_val,
                             "clocks": sustained[2]}
        frac["vs_sustained"] = s_val / peaks["sustained"]
    if strong:
        st_step = strong_ms / args.steps
        line["strong"] = {"scaling": "strong", "workload": "C3 with B=16 GLOBAL, batch-sharded over the ranks (configs[2] as worded)",
                          "global_batch": B, "ms_per_step": st_step, "value": (ff + fb) / (st_step * 1e-3) / 1e12,
                          "unit": "TFLOP/s", "per_gpu_tflops": (ff + fb) / (st_step * 1e-3) / 1e12 / world}
    if ring_c5:
        line["ring_c5"] = ring_c5
    if e2e:
        gbs = lambda nbytes: nbytes / (e2e_ms * 1e-3) / 1e9
        line["e2e"] = {"value": world * (ff + fb) / (e2e_ms * 1e-3) / 1e12, "unit": "TFLOP/s",
                       "h2d_bytes_per_step": e2e[1], "d2h_bytes_per_step": e2e[2], "ms_per_step": e2e_ms,
                       "how": f"pinned host buffers, {args.e2e_chunks} batch chunks pipelined on 3 streams (H2D | kernels | D2H; the forward starts once q, k, v are in, out leaves under the backward)",
                       "sequential_ms_per_step": e2e[3],
                       "pcie_gbs_per_gpu": {"h2d": gbs(e2e[1]), "d2h": gbs(e2e[2]),
                                            "note": "achieved per direction per GPU over the whole step; the step is "
                                                    "bound by these copies, not by the kernels"}}
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
    ap.add_argument("--no-sustained", action="store_true")
    ap.add_argument("--sustain-seconds", type=float, default=3.0)
    ap.add_argument("--warm-seconds", type=float, default=1.5)
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--no-ring", action="store_true")
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
