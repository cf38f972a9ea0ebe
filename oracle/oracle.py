"""ctypes binding of the CPU oracle (oracle/attn_oracle.c).  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module; the product package
(flash_attention_hip_b200) never does.  Parity is unpinned by the reference
(it ships no implementation or fixtures); see
attn_oracle.h for how the oracle is pinned instead.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libattn_oracle.so")


class _Params(ctypes.Structure):
    _fields_ = [
        ("batch", ctypes.c_int32),
        ("seqlen_q", ctypes.c_int32),
        ("seqlen_k", ctypes.c_int32),
        ("num_heads", ctypes.c_int32),
        ("num_heads_k", ctypes.c_int32),
        ("head_size", ctypes.c_int32),
        ("is_causal", ctypes.c_int32),
        ("is_bf16", ctypes.c_int32),
        ("use_fp64", ctypes.c_int32),
        ("seqlen_q_rounded", ctypes.c_int32),
        ("seqlen_k_rounded", ctypes.c_int32),
        ("p_dropout", ctypes.c_float),
        ("softmax_scale", ctypes.c_float),
        ("seed", ctypes.c_uint64),
        ("offset", ctypes.c_uint64),
        ("cu_seqlens_q", ctypes.c_void_p),
        ("cu_seqlens_k", ctypes.c_void_p),
        ("attn_mask", ctypes.c_void_p),
        ("mask_dims", ctypes.c_int64 * 4),
        ("batch_base", ctypes.c_int32),
        ("round_intermediates", ctypes.c_int32),
    ]


def build(force: bool = False) -> str:
    """Compile the oracle with gcc if the .so is missing (or ``force``)."""
    if force or not os.path.exists(_LIB_PATH):
        subprocess.run(["make", "-C", _HERE] + (["-B"] if force else []), check=True,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.fa_oracle_fwd.restype = ctypes.c_int
        _lib.fa_oracle_bwd.restype = ctypes.c_int
        _lib.fa_oracle_num_threads.restype = ctypes.c_int
        _lib.fa_oracle_f32_to_storage.restype = ctypes.c_uint16
        _lib.fa_oracle_f32_to_storage.argtypes = [ctypes.c_float, ctypes.c_int]
        _lib.fa_oracle_storage_to_f32.restype = ctypes.c_float
        _lib.fa_oracle_storage_to_f32.argtypes = [ctypes.c_uint16, ctypes.c_int]
    return _lib


def num_threads() -> int:
    return int(lib().fa_oracle_num_threads())


# ---- storage helpers (numpy has no bfloat16) --------------------------------

def to_storage(x: np.ndarray, is_bf16: bool) -> np.ndarray:
    """float32 array -> uint16 storage bits (round-to-nearest-even)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    if not is_bf16:
        return x.astype(np.float16).view(np.uint16)
    u = x.view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    nan = np.isnan(x)
    if nan.any():
        r[nan] = 0x7FC0
    return r


def from_storage(x: np.ndarray, is_bf16: bool) -> np.ndarray:
    """uint16 storage bits -> float32."""
    x = np.ascontiguousarray(x, dtype=np.uint16)
    if not is_bf16:
        return x.view(np.float16).astype(np.float32)
    return (x.astype(np.uint32) << 16).view(np.float32)


@dataclass
class Problem:
    """One attention problem in the layouts of reference flash_attn.h."""
    batch: int
    seqlen_q: int            # dense length, or max_seqlen_q when packed
    seqlen_k: int
    num_heads: int
    num_heads_k: int
    head_size: int
    is_causal: bool = False
    is_bf16: bool = True
    p_dropout: float = 0.0
    softmax_scale: Optional[float] = None
    seed: int = 0
    offset: int = 0
    cu_seqlens_q: Optional[np.ndarray] = None
    cu_seqlens_k: Optional[np.ndarray] = None
    attn_mask: Optional[np.ndarray] = None      # uint16 storage [mb, mh, sq, sk]
    use_fp64: bool = False
    batch_base: int = 0      # global index of batch 0 when this problem is a batch shard
    round_intermediates: bool = False   # the "16-bit pipeline" build: P and dS rounded to the storage dtype too

    def __post_init__(self):
        if self.softmax_scale is None:
            self.softmax_scale = float(self.head_size) ** -0.5
        if self.cu_seqlens_q is not None:
            self.cu_seqlens_q = np.ascontiguousarray(self.cu_seqlens_q, dtype=np.int32)
            self.cu_seqlens_k = np.ascontiguousarray(self.cu_seqlens_k, dtype=np.int32)

    @property
    def varlen(self) -> bool:
        return self.cu_seqlens_q is not None

    @property
    def total_q(self) -> int:
        return int(self.cu_seqlens_q[-1]) if self.varlen else self.batch * self.seqlen_q

    @property
    def total_k(self) -> int:
        return int(self.cu_seqlens_k[-1]) if self.varlen else self.batch * self.seqlen_k

    @property
    def seqlen_q_rounded(self) -> int:
        return (self.seqlen_q + 127) // 128 * 128

    @property
    def seqlen_k_rounded(self) -> int:
        return (self.seqlen_k + 127) // 128 * 128

    def q_shape(self):
        return (self.total_q, self.num_heads, self.head_size)

    def k_shape(self):
        return (self.total_k, self.num_heads_k, self.head_size)

    def _c(self) -> _Params:
        p = _Params()
        p.batch, p.seqlen_q, p.seqlen_k = self.batch, self.seqlen_q, self.seqlen_k
        p.num_heads, p.num_heads_k, p.head_size = self.num_heads, self.num_heads_k, self.head_size
        p.is_causal, p.is_bf16, p.use_fp64 = int(self.is_causal), int(self.is_bf16), int(self.use_fp64)
        p.seqlen_q_rounded, p.seqlen_k_rounded = self.seqlen_q_rounded, self.seqlen_k_rounded
        p.p_dropout, p.softmax_scale = self.p_dropout, self.softmax_scale
        p.seed, p.offset = self.seed, self.offset
        p.batch_base = self.batch_base
        p.round_intermediates = int(self.round_intermediates)
        if self.varlen:
            p.cu_seqlens_q = self.cu_seqlens_q.ctypes.data
            p.cu_seqlens_k = self.cu_seqlens_k.ctypes.data
        if self.attn_mask is not None:
            self.attn_mask = np.ascontiguousarray(self.attn_mask, dtype=np.uint16)
            p.attn_mask = self.attn_mask.ctypes.data
            for i in range(4):
                p.mask_dims[i] = self.attn_mask.shape[i]
        return p


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else ctypes.c_void_p(a.ctypes.data)


def forward(pb: Problem, q, k, v, return_softmax: bool = False):
    """q/k/v: uint16 storage arrays [total, heads, d].  Returns dict."""
    q, k, v = (np.ascontiguousarray(a, dtype=np.uint16) for a in (q, k, v))
    assert q.shape == pb.q_shape() and k.shape == pb.k_shape() and v.shape == pb.k_shape()
    out = np.zeros(pb.q_shape(), np.uint16)
    out_f32 = np.zeros(pb.q_shape(), np.float32)
    lse = np.full((pb.batch, pb.num_heads, pb.seqlen_q), np.inf, np.float32)
    sm = None
    if return_softmax:
        sm = np.zeros((pb.batch, pb.num_heads, pb.seqlen_q_rounded, pb.seqlen_k_rounded), np.uint16)
    cp = pb._c()
    rc = lib().fa_oracle_fwd(ctypes.byref(cp), _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(lse),
                             _ptr(out_f32), _ptr(sm))
    if rc:
        raise RuntimeError(f"fa_oracle_fwd failed rc={rc}")
    return {"out": out, "out_f32": out_f32, "lse": lse, "softmax": sm}


def backward(pb: Problem, dout, q, k, v, out, lse):
    dout, q, k, v, out = (np.ascontiguousarray(a, dtype=np.uint16) for a in (dout, q, k, v, out))
    lse = np.ascontiguousarray(lse, dtype=np.float32)
    dq = np.zeros(pb.q_shape(), np.uint16)
    dk = np.zeros(pb.k_shape(), np.uint16)
    dv = np.zeros(pb.k_shape(), np.uint16)
    dq32 = np.zeros(pb.q_shape(), np.float32)
    dk32 = np.zeros(pb.k_shape(), np.float32)
    dv32 = np.zeros(pb.k_shape(), np.float32)
    cp = pb._c()
    rc = lib().fa_oracle_bwd(ctypes.byref(cp), _ptr(dout), _ptr(q), _ptr(k), _ptr(v), _ptr(out),
                             _ptr(lse), _ptr(dq), _ptr(dk), _ptr(dv), _ptr(dq32), _ptr(dk32),
                             _ptr(dv32))
    if rc:
        raise RuntimeError(f"fa_oracle_bwd failed rc={rc}")
    return {"dq": dq, "dk": dk, "dv": dv, "dq_f32": dq32, "dk_f32": dk32, "dv_f32": dv32}


def dropout_mask(seed: int, offset: int, p_dropout: float, bh: int, sq: int, sk: int) -> np.ndarray:
    m = np.zeros((sq, sk), np.uint8)
    lib().fa_oracle_dropout_mask(ctypes.c_uint64(seed), ctypes.c_uint64(offset),
                                 ctypes.c_float(p_dropout), ctypes.c_uint32(bh),
                                 ctypes.c_int32(sq), ctypes.c_int32(sk), _ptr(m))
    return m


def philox(ctr, key, rounds: Optional[int] = None) -> np.ndarray:
    """Philox-4x32; rounds=None -> the dropout specification's round count (csrc/philox_spec.h)."""
    c = (ctypes.c_uint32 * 4)(*ctr)
    k = (ctypes.c_uint32 * 2)(*key)
    o = (ctypes.c_uint32 * 4)()
    if rounds is None:
        lib().fa_oracle_philox(c, k, o)
    else:
        lib().fa_oracle_philox_r(c, k, ctypes.c_int(rounds), o)
    return np.array(list(o), dtype=np.uint32)


def philox_rounds() -> int:
    return int(lib().fa_oracle_philox_rounds())
