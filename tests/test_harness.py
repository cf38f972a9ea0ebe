"""The Paddle-shaped C caller (tests/harness/paddle_like_caller.c): a plain-C
program that dlopen()s libflash_attn.so, binds the six reference names with dlsym and drives
fwd -> bwd with caller-allocated scratch.  CPU: it compiles as C against the CUDA runtime headers.
GPU: it runs, and its outputs match the oracle on the same LCG inputs (north_star tolerances)."""
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "harness", "paddle_like_caller.c")
EXE = os.path.join(ROOT, "tests", "harness", "paddle_like_caller")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def _build():
    subprocess.run(["gcc", "-std=c11", "-O2", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(CUDA, "include"),
                    SRC, "-o", EXE, "-L", os.path.join(CUDA, "lib64"), "-lcudart", "-ldl", "-lm",
                    "-Wl,-rpath," + os.path.join(CUDA, "lib64")], check=True)
    return EXE


def test_caller_compiles_as_plain_c():
    assert os.path.exists(_build())


def _lcg_bf16(n, state):
    """numpy restatement of next_bf16(): returns (bits uint16[n], new state)."""
    out = np.empty(n, np.uint16)
    # vectorised LCG: x_{i} = a^i x0 + c (a^i - 1)/(a - 1)  (mod 2^32), done iteratively in blocks
    a, c = np.uint64(1664525), np.uint64(1013904223)
    x = np.uint64(state)
    m = np.uint64(0xFFFFFFFF)
    xs = np.empty(n, np.uint64)
    for i in range(n):
        x = (x * a + c) & m
        xs[i] = x
    v = ((xs >> np.uint64(16)) & np.uint64(0xFF)).astype(np.int64) - 128
    f = (v / 128.0).astype(np.float32)
    out[:] = (f.view(np.uint32) >> 16).astype(np.uint16)
    return out, int(x)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(2, 256, 256, 4, 2, 128, 1, 0.1), (1, 200, 333, 4, 4, 64, 0, 0.0)])
def test_caller_runs_and_matches_oracle(shape, tmp_path):
    from flash_attention_hip_b200 import api
    from oracle import oracle as O
    exe = _build()
    b, sq, sk, h, hk, d, causal, p = shape
    outbin = str(tmp_path / "out.bin")
    r = subprocess.run([exe, api.LIB_PATH, outbin] + [str(x) for x in shape], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "expected failure:" in r.stdout and "ok: fwd+bwd" in r.stdout
    nq, nk = b * sq * h * d, b * sk * hk * d
    st = 12345
    q, st = _lcg_bf16(nq, st); k, st = _lcg_bf16(nk, st); v, st = _lcg_bf16(nk, st); do, st = _lcg_bf16(nq, st)
    pb = O.Problem(batch=b, seqlen_q=sq, seqlen_k=sk, num_heads=h, num_heads_k=hk, head_size=d,
                   is_causal=bool(causal), p_dropout=p, seed=1234, offset=0)
    ref = O.forward(pb, q.reshape(pb.q_shape()), k.reshape(pb.k_shape()), v.reshape(pb.k_shape()))
    rb = O.backward(pb, do.reshape(pb.q_shape()), q.reshape(pb.q_shape()), k.reshape(pb.k_shape()),
                    v.reshape(pb.k_shape()), ref["out"], ref["lse"])
    raw = open(outbin, "rb").read()
    hdr = struct.unpack("8i", raw[:32])
    assert hdr[:7] == (b, sq, sk, h, hk, d, causal) and hdr[7] == 1234   # rng_state[0] = seed
    off = 32
    def take(n, dt):
        nonlocal off
        a = np.frombuffer(raw, dt, n, off); off += n * np.dtype(dt).itemsize
        return a
    out, lse = take(nq, np.uint16), take(b * h * sq, np.float32)
    dq, dk, dv = take(nq, np.uint16), take(nk, np.uint16), take(nk, np.uint16)
    rel = lambda got, want: np.abs(O.from_storage(got, True) - want.ravel()).max() / np.abs(want).max()
    assert rel(out, ref["out_f32"]) < 2e-2
    assert np.abs(lse - ref["lse"].ravel()).max() < 1e-3
    assert rel(dq, rb["dq_f32"]) < 2e-2 and rel(dk, rb["dk_f32"]) < 2e-2 and rel(dv, rb["dv_f32"]) < 2e-2
