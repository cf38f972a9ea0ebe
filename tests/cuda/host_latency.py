"""Host-side cost of one entry-point call (validation, tensor-map encoding, launches) next to its device time, on the
shapes where it can matter: BASELINE.json config 1 and the short-sequence shapes.  The host time is the wall time of
N back-to-back calls without synchronisation divided by N, measured while the launch queue is NOT full (N small enough);
the device time the CUDA-event time of the same calls."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import ctypes
import torch
from flash_attention_hip_b200 import api


def case(b, s, h, hk, d, causal, n=64):
    q, do = (torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16) for _ in range(2))
    k, v = (torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16) for _ in range(2))
    out = torch.empty_like(q)
    lse = torch.empty((b, h, s), dtype=torch.float32, device="cuda")
    sq_r, d_r = api.round_up(s, 128), api.round_up(d, 32)
    scratch = (torch.empty((b, h, sq_r), dtype=torch.float32, device="cuda"), torch.empty((b, sq_r, h, d_r), dtype=torch.float32, device="cuda"))
    grads = (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    L = api.lib()
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    scale = d ** -0.5
    # straight through the C ABI (no Python wrapper work in the timed loop beyond the ctypes call itself)
    fwd = lambda: L.flash_attn_fwd(P(q), P(k), P(v), None, P(out), None, P(lse), b, s, s, sq_r, sq_r, h, hk, d, d_r, 0.0, scale, 1.0 / scale,
                                   causal, False, True, st, 0, 0, None, None)
    bwd = lambda: L.flash_attn_bwd(P(do), P(q), P(k), P(v), P(out), P(scratch[0]), P(lse), None, P(grads[0]), P(grads[1]), P(grads[2]),
                                   P(scratch[1]), b, s, s, sq_r, sq_r, h, hk, d, d_r, 0.0, scale, 1.0 / scale, causal, True, 0, st, 0, 0, None, None)
    for name, f in (("flash_attn_fwd", fwd), ("flash_attn_bwd", bwd)):
        for _ in range(8):
            assert f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(n):
            f()
        t1 = time.perf_counter()
        e1.record(); torch.cuda.synchronize()
        print(f"{name} b{b} s{s} h{h}/{hk} d{d} causal={causal}: host {1e6 * (t1 - t0) / n:7.1f} us per call, device {1e3 * e0.elapsed_time(e1) / n:8.1f} us per call "
              f"({api.last_launch_count()} launches)", flush=True)


case(1, 128, 4, 4, 64, False)          # config 1
case(256, 128, 16, 16, 128, False)
case(64, 256, 32, 32, 128, True)
case(8, 2048, 16, 16, 128, True)       # config 2
