"""Developer tool: run one fwd+bwd with the -DFA_TRACE build and print per-role event timelines
(cycles between consecutive events, averaged over the steady-state iterations).
  FA_B200_TRACE=1 python tests/cuda/trace_run.py [fwd|bwd] [p_dropout] [causal 0/1]"""
import ctypes, os, sys, collections
os.environ.setdefault("FA_B200_VARIANT", "trace")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from flash_attention_hip_b200 import api

which = sys.argv[1] if len(sys.argv) > 1 else "bwd"
p = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
causal = bool(int(sys.argv[3])) if len(sys.argv) > 3 else False
b, s, h, d = int(os.environ.get("TRACE_B", 2)), int(os.environ.get("TRACE_S", 4096)), int(os.environ.get("TRACE_H", 16)), 128
q, k, v, do = (torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16) for _ in range(4))
kw = dict(p_dropout=p, is_causal=causal, seed=1, offset=0)
out, lse, _, _ = api.flash_attn_fwd(q, k, v, **kw)
if which == "bwd":
    api.flash_attn_bwd(do, q, k, v, out, lse, **kw)
torch.cuda.synchronize()
lib = api.lib()
lib.flash_attn_b200_trace_read.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
buf = (ctypes.c_uint64 * 4096)()
for role in range(8):
    n = lib.flash_attn_b200_trace_read(0 if which == "fwd" else 1, role, buf, 4096)
    if n <= 0:
        continue
    ev = [(buf[i] >> 8, buf[i] & 0xff) for i in range(n)]
    wall = {e: t for t, e in ev if e in (254, 255)}
    clk = {e: t for t, e in ev if e in (252, 253)}
    ev = [x for x in ev if x[1] < 252]
    n = len(ev)
    if 254 in wall and 255 in wall and wall[255] > wall[254]:
        print(f"    SM clock during the traced CTA: {(clk[253] - clk[252]) / (wall[255] - wall[254]) * 1e3:.0f} MHz")
    t0 = ev[0][0]
    if os.environ.get("TRACE_RAW") == "1":
        base = int(os.environ.get("TRACE_T0", 0)) or t0
        if not os.environ.get("TRACE_T0"):
            os.environ["TRACE_T0"] = str(t0)
        print(f"--- role {role} raw: " + " ".join(f"{e}@{t - base}" for t, e in ev[:int(os.environ.get('TRACE_N', 60))]))
        continue
    print(f"--- role {role}: {n} events, span {ev[-1][0] - t0} cycles")
    # mean delta between consecutive event ids over the middle half
    lo, hi = n // 4, 3 * n // 4
    agg = collections.defaultdict(list)
    for i in range(max(lo, 1), hi):
        agg[(ev[i - 1][1], ev[i][1])].append(ev[i][0] - ev[i - 1][0])
    for (a, b_), vals in sorted(agg.items()):
        print(f"   {a:3d} -> {b_:3d}: mean {sum(vals) / len(vals):8.0f}  min {min(vals):6d} max {max(vals):6d}  n={len(vals)}")
    for i in range(1, min(n, 40)):
        if which == "fwd" and (ev[i][1] >= 20 or ev[i - 1][1] >= 20):
            print(f"   one-off {ev[i - 1][1]:3d} -> {ev[i][1]:3d}: {ev[i][0] - ev[i - 1][0]}")
    first = ev[0][1]
    idx = [t for t, e in ev if e == first]
    if len(idx) > 4:
        per = [idx[i + 1] - idx[i] for i in range(len(idx) // 4, 3 * len(idx) // 4)]
        print(f"   period of event {first}: mean {sum(per) / len(per):.0f} cycles")
