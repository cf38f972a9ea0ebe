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


def run_fwd(pb: O.Problem, q, k, v, return_softmax=False):
    """Returns dict(out bits, lse, softmax bits|None, rng_state) from the CUDA library."""
    dq, dk, dv = (to_dev(a, pb.is_bf16) for a in (q, k, v))
    kw = dict(p_dropout=pb.p_dropout, softmax_scale=pb.softmax_scale, is_causal=pb.is_causal,
              return_softmax=return_softmax, seed=pb.seed, offset=pb.offset)
    if pb.varlen:
        cq = torch.from_numpy(pb.cu_seqlens_q).cuda()
        ck = torch.from_numpy(pb.cu_seqlens_k).cuda()
        out, lse, sm, rng = api.flash_attn_varlen_fwd(dq, dk, dv, cq, ck, pb.seqlen_q, pb.seqlen_k, **kw)
    else:
        shp_q = (pb.batch, pb.seqlen_q, pb.num_heads, pb.head_size)
        shp_k = (pb.batch, pb.seqlen_k, pb.num_heads_k, pb.head_size)
        mask = None if pb.attn_mask is None else to_dev(pb.attn_mask, pb.is_bf16)
        out, lse, sm, rng = api.flash_attn_fwd(dq.view(shp_q), dk.view(shp_k), dv.view(shp_k),
                                               attn_mask=mask, **kw)
    torch.cuda.synchronize()
    return dict(out=to_bits(out).reshape(pb.q_shape()), lse=lse.cpu().numpy(),
                softmax=None if sm is None else to_bits(sm), rng=rng.cpu().numpy(), path=api.last_path())


def run_bwd(pb: O.Problem, dout, q, k, v, out, lse, deterministic=False):
    t = [to_dev(a, pb.is_bf16) for a in (dout, q, k, v, out)]
    lse_d = torch.from_numpy(np.ascontiguousarray(lse, dtype=np.float32)).cuda()
    kw = dict(p_dropout=pb.p_dropout, softmax_scale=pb.softmax_scale, is_causal=pb.is_causal,
              deterministic=deterministic, seed=pb.seed, offset=pb.offset)
    if pb.varlen:
        cq = torch.from_numpy(pb.cu_seqlens_q).cuda()
        ck = torch.from_numpy(pb.cu_seqlens_k).cuda()
        dq, dk, dv = api.flash_attn_varlen_bwd(*t, lse_d, cq, ck, pb.seqlen_q, pb.seqlen_k, **kw)
    else:
        shp_q = (pb.batch, pb.seqlen_q, pb.num_heads, pb.head_size)
        shp_k = (pb.batch, pb.seqlen_k, pb.num_heads_k, pb.head_size)
        mask = None if pb.attn_mask is None else to_dev(pb.attn_mask, pb.is_bf16)
        dq, dk, dv = api.flash_attn_bwd(t[0].view(shp_q), t[1].view(shp_q), t[2].view(shp_k),
                                        t[3].view(shp_k), t[4].view(shp_q), lse_d, attn_mask=mask, **kw)
    torch.cuda.synchronize()
    return dict(dq=to_bits(dq).reshape(pb.q_shape()), dk=to_bits(dk).reshape(pb.k_shape()),
                dv=to_bits(dv).reshape(pb.k_shape()), path=api.last_path())


def f32(bits, pb):
    return O.from_storage(bits, pb.is_bf16)
