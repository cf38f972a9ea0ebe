"""Helpers for the -m gpu parity tests: numpy storage bits <-> CUDA tensors, and
running one oracle Problem through the C ABI (flash_attention_hip_b200.api)."""
import numpy as np
import torch

from flash_attention_hip_b200 import api
from oracle import oracle as O


def to_dev(bits: np.ndarray, is_bf16: bool) -> torch.Tensor:
    t = torch.from_numpy(np.ascontiguousarray(bits).view(np.int16))
    return t.view(torch.bfloat16 if is_bf16 else torch.float16).cuda()


def to_bits(t: torch.Tensor) -> np.ndarray:
    return t.detach().cpu().contiguous().view(torch.int16).numpy().view(np.uint16)


class Guarded:
    """Output buffers with 4 KiB of a known byte pattern on both sides: every run_fwd / run_bwd
    checks that the library wrote nothing outside the sizes the header implies (scratch included:
    softmax_d fp32 [b, h, sq_rounded], dq_accum fp32 b * sq_rounded * h * d_rounded, no more)."""
    PAD, BYTE = 4096, 0x5A

    def __init__(self):
        self.bufs = []

    def make(self, shape, dtype, fill=None):
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        raw = torch.full((n + 2 * self.PAD,), self.BYTE, dtype=torch.uint8, device="cuda")
        self.bufs.append((raw, n))
        t = raw[self.PAD:self.PAD + n].view(dtype).view(shape)
        if fill is not None:
            t.fill_(fill)
        return t

    def check(self):
        for raw, n in self.bufs:
            assert bool((raw[:self.PAD] == self.BYTE).all()), "write below an output buffer"
            assert bool((raw[self.PAD + n:] == self.BYTE).all()), "write past an output buffer"


def oracle_fwd(pb: O.Problem, q, k, v):
    """(exact oracle forward, the 16-bit-pipeline build's out as f32): the reference and the error baseline of
    tests/parity.py's second criterion."""
    import dataclasses
    ref = O.forward(pb, q, k, v)
    lp = O.forward(dataclasses.replace(pb, round_intermediates=True), q, k, v)
    return ref, O.from_storage(lp["out"], pb.is_bf16)


def oracle_bwd(pb: O.Problem, dout, q, k, v, out, lse):
    """(exact oracle backward, {dq, dk, dv} of the 16-bit-pipeline build as f32)."""
    import dataclasses
    rb = O.backward(pb, dout, q, k, v, out, lse)
    lp = O.backward(dataclasses.replace(pb, round_intermediates=True), dout, q, k, v, out, lse)
    return rb, {t: O.from_storage(lp[t], pb.is_bf16) for t in ("dq", "dk", "dv")}


def check_fwd(label, pb, got, ref, lp_out, varlen_lse=False):
    """All parity criteria (tests/parity.py) on one forward result."""
    import parity
    import cases
    parity.check(label, "out", f32(got["out"], pb), ref["out_f32"], lp_out)
    if pb.varlen or varlen_lse:      # rows past a sequence's end are unspecified
        for b, _, sq, _, _ in cases.seq_slices(pb):
            parity.check_lse(f"{label}[seq {b}]", got["lse"][b, :, :sq], ref["lse"][b, :, :sq])
    else:
        parity.check_lse(label, got["lse"], ref["lse"])


def check_bwd(label, pb, gb, rb, lp):
    import parity
    for t in ("dq", "dk", "dv"):
        parity.check(label, t, f32(gb[t], pb), rb[t + "_f32"], lp[t] if lp is not None else None)


def run_fwd(pb: O.Problem, q, k, v, return_softmax=False):
    """Returns dict(out bits, lse, softmax bits|None, rng_state) from the CUDA library."""
    dq, dk, dv = (to_dev(a, pb.is_bf16) for a in (q, k, v))
    kw = dict(p_dropout=pb.p_dropout, softmax_scale=pb.softmax_scale, is_causal=pb.is_causal,
              return_softmax=return_softmax, seed=pb.seed, offset=pb.offset, batch_base=pb.batch_base)
    g = Guarded()
    nb = len(pb.cu_seqlens_q) - 1 if pb.varlen else pb.batch
    kw["lse"] = g.make((nb, pb.num_heads, pb.seqlen_q), torch.float32, fill=float("inf"))
    if pb.varlen:
        cq = torch.from_numpy(pb.cu_seqlens_q).cuda()
        ck = torch.from_numpy(pb.cu_seqlens_k).cuda()
        kw["out"] = g.make(tuple(dq.shape), dq.dtype)
        mask = None if pb.attn_mask is None else to_dev(pb.attn_mask, pb.is_bf16)
        out, lse, sm, rng = api.flash_attn_varlen_fwd(dq, dk, dv, cq, ck, pb.seqlen_q, pb.seqlen_k, attn_mask=mask, **kw)
    else:
        shp_q = (pb.batch, pb.seqlen_q, pb.num_heads, pb.head_size)
        shp_k = (pb.batch, pb.seqlen_k, pb.num_heads_k, pb.head_size)
        mask = None if pb.attn_mask is None else to_dev(pb.attn_mask, pb.is_bf16)
        kw["out"] = g.make(shp_q, dq.dtype)
        out, lse, sm, rng = api.flash_attn_fwd(dq.view(shp_q), dk.view(shp_k), dv.view(shp_k),
                                               attn_mask=mask, **kw)
    torch.cuda.synchronize()
    g.check()
    return dict(out=to_bits(out).reshape(pb.q_shape()), lse=lse.cpu().numpy(),
                softmax=None if sm is None else to_bits(sm), rng=rng.cpu().numpy(), path=api.last_path())


def run_bwd(pb: O.Problem, dout, q, k, v, out, lse, deterministic=False):
    t = [to_dev(a, pb.is_bf16) for a in (dout, q, k, v, out)]
    lse_d = torch.from_numpy(np.ascontiguousarray(lse, dtype=np.float32)).cuda()
    kw = dict(p_dropout=pb.p_dropout, softmax_scale=pb.softmax_scale, is_causal=pb.is_causal,
              deterministic=deterministic, seed=pb.seed, offset=pb.offset, batch_base=pb.batch_base)
    g = Guarded()
    nb = len(pb.cu_seqlens_q) - 1 if pb.varlen else pb.batch
    sq_r, d_r = api.round_up(pb.seqlen_q, 128), api.round_up(pb.head_size, 32)
    kw["scratch"] = (g.make((nb, pb.num_heads, sq_r), torch.float32),
                     g.make((nb, sq_r, pb.num_heads, d_r), torch.float32))
    if pb.varlen:
        cq = torch.from_numpy(pb.cu_seqlens_q).cuda()
        ck = torch.from_numpy(pb.cu_seqlens_k).cuda()
        kw["grads"] = (g.make(tuple(t[1].shape), t[1].dtype), g.make(tuple(t[2].shape), t[2].dtype),
                       g.make(tuple(t[3].shape), t[3].dtype))
        mask = None if pb.attn_mask is None else to_dev(pb.attn_mask, pb.is_bf16)
        dq, dk, dv = api.flash_attn_varlen_bwd(*t, lse_d, cq, ck, pb.seqlen_q, pb.seqlen_k, attn_mask=mask, **kw)
    else:
        shp_q = (pb.batch, pb.seqlen_q, pb.num_heads, pb.head_size)
        shp_k = (pb.batch, pb.seqlen_k, pb.num_heads_k, pb.head_size)
        mask = None if pb.attn_mask is None else to_dev(pb.attn_mask, pb.is_bf16)
        kw["grads"] = (g.make(shp_q, t[1].dtype), g.make(shp_k, t[2].dtype), g.make(shp_k, t[3].dtype))
        dq, dk, dv = api.flash_attn_bwd(t[0].view(shp_q), t[1].view(shp_q), t[2].view(shp_k),
                                        t[3].view(shp_k), t[4].view(shp_q), lse_d, attn_mask=mask, **kw)
    torch.cuda.synchronize()
    g.check()
    return dict(dq=to_bits(dq).reshape(pb.q_shape()), dk=to_bits(dk).reshape(pb.k_shape()),
                dv=to_bits(dv).reshape(pb.k_shape()), path=api.last_path())


def f32(bits, pb):
    return O.from_storage(bits, pb.is_bf16)
