"""Independent restatement of ring context-parallel attention WITH dropout (include/flash_attn_b200_ext.h, "Dropout"):
full-sequence attention in float64 numpy under a keep mask ASSEMBLED from the per-step masks of the written
specification -- step (rank r, K/V owner src) draws csrc/philox_spec.h on the rows and columns of its own
(query slice, K/V chunk) with offset + ((r * N + src) << 48) and head word (batch entry) * num_heads + head.
Nothing here calls the ring schedule under test; only its slice table (`ring._steps`) is shared."""
import numpy as np

import ref_numpy


def positions(rank, world, s_local, causal):
    """Global sequence positions of a rank's local rows (zig-zag when causal, contiguous otherwise)."""
    if not causal:
        return np.arange(rank * s_local, (rank + 1) * s_local)
    c = s_local // 2
    return np.concatenate([np.arange(rank * c, (rank + 1) * c), np.arange((2 * world - 1 - rank) * c, (2 * world - rank) * c)])


def composite_keep(world, s_local, causal, p_dropout, seed, offset, bh):
    """[S, S] bool: the keep bit every (query, key) pair gets from the one ring step that computes it."""
    from flash_attention_hip_b200 import ring
    S = world * s_local
    keep = np.ones((S, S), bool)
    for r in range(world):
        for src in range(world):
            (q0, q1), (k0, k1), _ = ring._steps(causal, r, src, s_local)
            m = ref_numpy.keep_mask(seed, ring.step_offset(offset, r, src, world), p_dropout, bh, q1 - q0, k1 - k0)
            rows = positions(r, world, s_local, causal)[q0:q1]
            cols = positions(src, world, s_local, causal)[k0:k1]
            keep[np.ix_(rows, cols)] = m
    return keep


def attention_with_keep(q, k, v, do, scale, causal, keep, p_dropout):
    """One (batch, head) in float64: q, do [S, d]; k, v [S, d]; keep [S, S].  Returns out, lse, dq, dk, dv."""
    S = q.shape[0]
    s = scale * (q @ k.T)
    if causal:
        s = np.where(np.arange(S)[None, :] <= np.arange(S)[:, None], s, -np.inf)
    m = s.max(axis=1, keepdims=True)
    e = np.exp(s - m)
    l = e.sum(axis=1, keepdims=True)
    p = e / l
    lse = (m + np.log(l))[:, 0]
    rp = 1.0 / (1.0 - p_dropout) if p_dropout < 1.0 else 0.0
    pd = np.where(keep, p * rp, 0.0)
    out = pd @ v
    dv = pd.T @ do
    dp = np.where(keep, (do @ v.T) * rp, 0.0)
    D = (do * out).sum(axis=1, keepdims=True)
    ds = p * (dp - D)
    return out, lse, scale * ds @ k, scale * ds.T @ q, dv


def full_reference(q, k, v, do, scale, causal, world, p_dropout, seed, offset):
    """q, do [b, S, h, d]; k, v [b, S, hk, d] float arrays.  Returns dict of float64 arrays in the same layouts
    (lse [b, h, S])."""
    b, S, h, d = q.shape
    hk = k.shape[2]
    g = h // hk
    s_local = S // world
    out = np.zeros((b, S, h, d)); dq = np.zeros((b, S, h, d))
    dk = np.zeros((b, S, hk, d)); dv = np.zeros((b, S, hk, d))
    lse = np.zeros((b, h, S))
    for bi in range(b):
        for hh in range(h):
            keep = composite_keep(world, s_local, causal, p_dropout, seed, offset, bi * h + hh) if p_dropout > 0 \
                else np.ones((S, S), bool)
            o, l, gq, gk, gv = attention_with_keep(q[bi, :, hh].astype(np.float64), k[bi, :, hh // g].astype(np.float64),
                                                   v[bi, :, hh // g].astype(np.float64), do[bi, :, hh].astype(np.float64),
                                                   scale, causal, keep, p_dropout)
            out[bi, :, hh] = o; lse[bi, hh] = l; dq[bi, :, hh] = gq
            dk[bi, :, hh // g] += gk; dv[bi, :, hh // g] += gv
    return dict(out=out, lse=lse, dq=dq, dk=dk, dv=dv)
