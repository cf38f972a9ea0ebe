#!/usr/bin/env python
"""Top stall sites of one kernel from an .ncu-rep source page (SASS view).
  python profiles/top_stalls.py REP KERNEL_REGEX [N]"""
import csv, io, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + rx],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
# possibly several kernels; take the first block
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[2:]:
    if len(r) < len(hdr) or r[0] == "Address":
        break
    body.append(r)
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[idx["# Samples"]]) for r in body)
print(f"# {rows[0][1][:100]}  total samples {tot}, {len(body)} SASS instructions")
agg = {s: sum(int(r[idx[s]]) for r in body) for s in stalls}
print("# by reason:", ", ".join(f"{k[6:]}={v} ({100*v/tot:.1f}%)" for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v))
order = sorted(range(len(body)), key=lambda i: -int(body[i][idx["# Samples"]]))[:n]
for i in sorted(order):
    r = body[i]
    why = sorted(((int(r[idx[s]]), s[6:]) for s in stalls), reverse=True)[:2]
    print(f"{i:5d} {int(r[idx['# Samples']]):7d} {100*int(r[idx['# Samples']])/tot:5.1f}%  {r[idx['Source']].strip()[:70]:70s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}")
