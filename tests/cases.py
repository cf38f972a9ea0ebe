"""Seeded synthetic problems shared by the CPU and GPU test suites."""
import numpy as np

from oracle import oracle as O


def make_inputs(pb: "O.Problem", seed: int, with_dout: bool = True, std: float = 1.0):
    """N(0,1) q/k/v (and dout) rounded to the storage dtype, as uint16 bits."""
    rng = np.random.default_rng(seed)
    q = O.to_storage(rng.standard_normal(pb.q_shape(), dtype=np.float32) * std, pb.is_bf16)
    k = O.to_storage(rng.standard_normal(pb.k_shape(), dtype=np.float32) * std, pb.is_bf16)
    v = O.to_storage(rng.standard_normal(pb.k_shape(), dtype=np.float32) * std, pb.is_bf16)
    dout = None
    if with_dout:
        dout = O.to_storage(rng.standard_normal(pb.q_shape(), dtype=np.float32) * std, pb.is_bf16)
    return q, k, v, dout


def cu_from_lens(lens):
    return np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)


def dense(b, sq, sk, h, hk, d, **kw):
    return O.Problem(batch=b, seqlen_q=sq, seqlen_k=sk, num_heads=h, num_heads_k=hk, head_size=d, **kw)


def varlen(lens_q, lens_k, h, hk, d, **kw):
    return O.Problem(batch=len(lens_q), seqlen_q=int(max(lens_q)), seqlen_k=int(max(lens_k)),
                     num_heads=h, num_heads_k=hk, head_size=d,
                     cu_seqlens_q=cu_from_lens(lens_q), cu_seqlens_k=cu_from_lens(lens_k), **kw)


def seq_slices(pb):
    """Yield (b, q0, sq, k0, sk) for every sequence of the problem."""
    for b in range(pb.batch):
        if pb.varlen:
            q0, q1 = int(pb.cu_seqlens_q[b]), int(pb.cu_seqlens_q[b + 1])
            k0, k1 = int(pb.cu_seqlens_k[b]), int(pb.cu_seqlens_k[b + 1])
        else:
            q0, q1 = b * pb.seqlen_q, (b + 1) * pb.seqlen_q
            k0, k1 = b * pb.seqlen_k, (b + 1) * pb.seqlen_k
        yield b, q0, q1 - q0, k0, k1 - k0


def rel_err(a, ref):
    """max |a-ref| / max(|ref|) -- the 'relative' of north_star's 2e-2 tolerance."""
    a = np.asarray(a, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    den = np.abs(ref).max()
    if den == 0:
        return float(np.abs(a).max())
    return float(np.abs(a - ref).max() / den)
