"""Writes the per-kernel SASS evidence file: for every kernel in libflash_attn.so, the counts of the
mnemonics that prove the Blackwell path (B200_PROFILING.md: UTC*MMA = tcgen05.mma, LDTM/STTM =
tcgen05.ld/st, UTMALDG/UTMAREDG/UBLKCP/UBLKRED = TMA / bulk copies, HMMA = legacy mma.sync) plus
registers and local-memory traffic.   python profiles/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "flash_attention_hip_b200", "lib", "libflash_attn.so")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
KEYS = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UBLKRED", "UTCBAR", "SYNCS", "MUFU.EX2",
        "FFMA2", "FMNMX3", "HMMA", "STL", "LDL", "USETMAXREG", "SHFL", "IMAD.WIDE"]
cur, counts, total = None, collections.OrderedDict(), collections.Counter()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = cur.replace("(anonymous namespace)::", "").replace("void ", "")
        cur = re.sub(r"\(.*", "", cur)
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        total[cur] += 1
        for k in KEYS:
            if op == k or op.startswith(k + ".") or (k in ("UTCHMMA",) and op.startswith(k)):
                counts[cur][k] += 1
print(f"# SASS mnemonic counts per kernel, {os.path.relpath(LIB, ROOT)} (cuobjdump -sass), sm_100a")
for name, c in counts.items():
    print(f"\n{name}\n  instructions {total[name]}")
    print("  " + "  ".join(f"{k}={c[k]}" for k in KEYS if c[k]))
