"""Builds libflash_attn.so (the C-ABI library) in-tree with nvcc for sm_100a.

`python -m flash_attention_hip_b200.build [--force] [--verbose]`
The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "libflash_attn.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
         "-Xptxas", "-v", "-ccbin", "/usr/bin/g++"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_digest() -> str:
    h = hashlib.sha1()
    for root in (CSRC, os.path.join(os.path.dirname(HERE), "include")):
        for f in sorted(os.listdir(root)):
            if f.endswith((".h", ".cuh", ".inc")):
                with open(os.path.join(root, f), "rb") as fh:
                    h.update(fh.read())
    h.update(" ".join(FLAGS + ARCH).encode())
    return h.hexdigest()


def _compile_one(src: str, digest: str, force: bool, verbose: bool, extra=(), objdir=None) -> str:
    objdir = objdir or OBJDIR
    digest = digest + " ".join(extra)
    obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
    stamp = obj + ".stamp"
    with open(src, "rb") as fh:
        key = hashlib.sha1(fh.read() + digest.encode()).hexdigest()
    if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == key:
        return obj
    cmd = [NVCC] + ARCH + FLAGS + list(extra) + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log = os.path.join(objdir, os.path.basename(src)[:-3] + ".log")
    with open(log, "w") as fh:
        fh.write(r.stdout)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout[-6000:]}")
    if verbose:
        print(r.stdout)
    with open(stamp, "w") as fh:
        fh.write(key)
    return obj


def build(force: bool = False, verbose: bool = False, trace: bool = False, variant: str = "", defs=()) -> str:
    """trace=True builds the developer variant libflash_attn_trace.so (-DFA_TRACE, csrc/fa_trace.cuh).
    variant="x", defs=("-DFOO=1",) builds a tuning variant libflash_attn_x.so (api.py: FA_B200_VARIANT=x)."""
    if trace:
        variant, defs = variant or "trace", tuple(defs) + ("-DFA_TRACE",)
    objdir = OBJDIR + ("_" + variant if variant else "")
    lib = LIB.replace(".so", "_" + variant + ".so") if variant else LIB
    extra = tuple(defs)
    os.makedirs(objdir, exist_ok=True)
    digest = _headers_digest()
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, digest, force, verbose, extra, objdir), srcs))
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(lib) or os.path.getmtime(lib) < newest:
        cmd = [NVCC] + ARCH + ["-shared", "-o", lib] + objs + ["-cudart", "static", "-ccbin", "/usr/bin/g++",
                                                      "-Xlinker", "--version-script=" + os.path.join(CSRC, "exports.map")]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout[-4000:]}")
    return lib


if __name__ == "__main__":
    var = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--variant=")]
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, trace="--trace" in sys.argv,
                 variant=var[0] if var else "", defs=[a for a in sys.argv if a.startswith("-D")])
    print(path)
