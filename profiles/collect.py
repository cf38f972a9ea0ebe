#!/usr/bin/env python
"""Copies one GPU round's outputs from gpurun_out/ (scratch) into profiles/ under a tag, and rebuilds the
derived evidence: ncu summary, DRAM traffic per launch (roofline_traffic.json), SASS mnemonic counts, full
SASS of the two D128 bf16 dropout kernels, ptxas -v summary.
  python profiles/collect.py r01e        (after tests/cuda/gpu_round.sh and gpu_ncu_full.sh ran on the box)"""
import json, os, re, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
for src, dst in (("bench.log", "bench.json"), ("bench_ref.log", "bench_reference.json"), ("launches.csv", "bench_launches.csv"),
                 ("quick_bench.log", "quick_bench.log"), ("red_probe.log", "red_probe.log"), ("bench_n2.log", "bench_n2.json")):
    if os.path.exists(os.path.join(G, src)):
        shutil.copy(os.path.join(G, src), os.path.join(P, f"{tag}_{dst}"))
tl = [os.path.join(G, f) for f in ("trace_fwd.log", "trace_fwd_drop.log", "trace_bwd.log", "trace_bwd_drop.log")]
if all(os.path.exists(f) for f in tl):
    with open(os.path.join(P, f"{tag}_fwd_timelines.log"), "w") as fh:
        fh.write("=== forward, no dropout\n" + open(tl[0]).read() + "=== forward, dropout 0.1\n" + open(tl[1]).read())
    with open(os.path.join(P, f"{tag}_bwd_timelines.log"), "w") as fh:
        fh.write("=== backward, no dropout\n" + open(tl[2]).read() + "=== backward, dropout 0.1\n" + open(tl[3]).read())
rep = os.path.join(G, "prof_fwd_bwd.ncu-rep")
pre = os.path.join(G, "c3_fwd_bwd_ncu.txt")          # summarised on the GPU box already (the reports exceed gpurun's copy-back limit)
for src, dst in (("wide_bench.log", "wide_bwd_bench.log"), ("det_bench.log", "det_bench.log"), ("wide_det_ncu.txt", "wide_det_ncu.txt"),
                 ("c3_fwd_bwd_stalls.txt", "c3_fwd_bwd_stalls.txt"), ("parity_errors.log", "parity_errors.log"),
                 ("short_end_launches.csv", "short_end_launches.csv"), ("host_latency.log", "host_latency.log")):
    if os.path.exists(os.path.join(G, src)):
        shutil.copy(os.path.join(G, src), os.path.join(P, f"{tag}_{dst}"))
if os.path.exists(rep) or os.path.exists(pre):
    if os.path.exists(rep):
        txt = subprocess.run([sys.executable, os.path.join(P, "summarize.py"), rep], capture_output=True, text=True).stdout
    else:
        txt = open(pre).read()
    open(os.path.join(P, f"{tag}_c3_fwd_bwd_ncu.txt"), "w").write(txt)
    traffic, cur = {}, None
    for line in txt.splitlines():
        m = re.match(r"== .*?(tc_fwd_kernel|tc_bwd_kernel|bwd_preprocess\w*|dq_convert_kernel)", line)
        if m:
            cur = m.group(1) if m.group(1) in traffic or True else None
            cur = "bwd_preprocess_kernel" if cur.startswith("bwd_preprocess") else cur
            if cur in traffic:
                cur = None          # first captured launch of each kernel only
            else:
                traffic[cur] = 0.0
            continue
        m = re.match(r"\s+dram__bytes_(read|write)\.sum\s+([0-9.]+) (G|M)byte", line)
        if m and cur:
            traffic[cur] += float(m.group(2)) * (1e9 if m.group(3) == "G" else 1e6)
    path = os.path.join(P, "roofline_traffic.json")
    old = json.load(open(path))
    old["_comment"] = (f"dram__bytes_read.sum + dram__bytes_write.sum per launch from ONE ncu --set full capture of the bench "
                       f"workload (C3: B16 H32 S4096 D128 causal dropout 0.1), profiles/{tag}_c3_fwd_bwd_ncu.txt; bench.py "
                       f"copies tc_bwd_kernel's figure into roofline.traffic")
    for k, v in traffic.items():
        old.setdefault(k, {})["dram_bytes_per_launch"] = int(round(v, -3))
    json.dump(old, open(path, "w"), indent=1)
    print("traffic", {k: round(v / 1e9, 3) for k, v in traffic.items()})
txt = subprocess.run([sys.executable, os.path.join(P, "sass_summary.py")], capture_output=True, text=True).stdout
open(os.path.join(P, f"{tag}_sass_summary.txt"), "w").write(txt)
objdir = os.path.join(ROOT, "flash_attention_hip_b200", "lib", "obj")
for name, obj, pat in (("tc_fwd_d128_bf16_dropout", "fa_tc_fwd.o", "tc_fwd_kernelILi128ELb1ELb1ELb0"),
                       ("tc_bwd_d128_bf16_dropout", "fa_tc_bwd.o", "tc_bwd_kernelILi128ELb1ELb1ELb0ELi0")):
    sass = subprocess.run(["cuobjdump", "-sass", os.path.join(objdir, obj)], capture_output=True, text=True).stdout
    out, on = [], False
    for line in sass.splitlines():
        if "Function :" in line:
            on = pat in line
        if on:
            out.append(line)
    open(os.path.join(P, f"{tag}_{name}.sass"), "w").write("\n".join(out) + "\n")
lines = []
for f in sorted(os.listdir(objdir)):
    if f.endswith(".log"):
        cur = None
        for line in open(os.path.join(objdir, f)):
            m = re.search(r"Compiling entry function '(\S+)'", line)
            if m:
                m2 = re.search(r"\d+((?:tc_|bwd_|dq_|generic_|mma_|cp_|write_)\w*?kernel\w*?)(?:EEv|EPm|EvP|$)", m.group(1))
                cur = m2.group(1) if m2 else m.group(1)
                lines.append(cur)
            elif cur and ("spill" in line or "Used" in line):
                lines.append(line.rstrip())
open(os.path.join(P, f"{tag}_ptxas_v.txt"), "w").write("\n".join(lines) + "\n")
print("collected", tag)
