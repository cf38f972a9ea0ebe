"""Independent numpy float64 restatement of the attention semantics.

Written separately from oracle/attn_oracle.c (no shared code, its own Philox)
so that agreement between the two pins the oracle.  Follows the layouts of
/root/reference/flash_attn.h:10-14 (dense), :38-44 (packed) and the decisions
in DESIGN.md section 3 / csrc/philox_spec.h.
"""
import numpy as np

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK32 = 0xFFFFFFFF


ROUNDS = 7          # csrc/philox_spec.h (restated here, not imported: this file is independent of it)


def philox4x32(ctr, key, rounds=ROUNDS):
    c0, c1, c2, c3 = (int(x) for x in ctr)
    k0, k1 = (int(x) for x in key)
    for _ in range(rounds):
        p0 = M0 * c0
        p1 = M1 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & MASK32, p1 & MASK32, \
                         ((p0 >> 32) ^ c3 ^ k1) & MASK32, p0 & MASK32
        k0 = (k0 + W0) & MASK32
        k1 = (k1 + W1) & MASK32
    return [c0, c1, c2, c3]


def philox4x32_vec(c0, c1, c2, c3, k0, k1, rounds=ROUNDS):
    """Vectorised over numpy uint64 arrays holding 32-bit values."""
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint64) for x in (c0, c1, c2, c3))
    k0 = np.uint64(k0)
    k1 = np.uint64(k1)
    m = np.uint64(MASK32)
    for _ in range(rounds):
        p0 = np.uint64(M0) * c0
        p1 = np.uint64(M1) * c2
        n0 = ((p1 >> np.uint64(32)) ^ c1 ^ k0) & m
        n1 = p1 & m
        n2 = ((p0 >> np.uint64(32)) ^ c3 ^ k1) & m
        n3 = p0 & m
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + np.uint64(W0)) & m
        k1 = (k1 + np.uint64(W1)) & m
    return c0, c1, c2, c3


def keep_mask(seed, offset, p_dropout, bh, sq, sk):
    """[sq, sk] bool, True = keep.  Spec: csrc/philox_spec.h."""
    thr = int(np.floor((1.0 - np.float32(p_dropout)) * np.float32(256.0) + np.float32(0.5)))
    thr = max(0, min(256, thr))
    nblk = (sk + 15) // 16
    rows = np.repeat(np.arange(sq, dtype=np.uint64), nblk)
    blks = np.tile(np.arange(nblk, dtype=np.uint64), sq)
    k0 = (seed & MASK32) ^ ((offset >> 32) & MASK32)
    k1 = (seed >> 32) & MASK32
    w = philox4x32_vec(rows, blks, np.full_like(rows, bh), np.full_like(rows, offset & MASK32),
                          k0, k1)
    by = np.zeros((sq * nblk, 16), dtype=np.uint64)
    for i in range(4):
        for j in range(4):
            by[:, 4 * i + j] = (w[i] >> np.uint64(8 * j)) & np.uint64(0xFF)
    by = by.reshape(sq, nblk * 16)[:, :sk]
    return by < thr


def attention_fwd(q, k, v, scale, causal=False, p_dropout=0.0, seed=0, offset=0, bh=0, bias=None):
    """Single (b, h): q [sq, d], k/v [sk, d] float64.  Returns out, lse, P, keep."""
    sq, sk = q.shape[0], k.shape[0]
    s = scale * (q @ k.T)
    if bias is not None:
        s = s + bias
    if causal:
        i = np.arange(sq)[:, None]
        j = np.arange(sk)[None, :]
        s = np.where(j <= i + sk - sq, s, -np.inf)
    m = s.max(axis=1, keepdims=True) if sk > 0 else np.full((sq, 1), -np.inf)
    msafe = np.where(np.isfinite(m), m, 0.0)
    e = np.exp(s - msafe)
    l = e.sum(axis=1, keepdims=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        lse = np.where(l > 0, msafe + np.log(l), np.inf)[:, 0]
        p = np.where(l > 0, e / l, 0.0)
    keep = np.ones((sq, sk), bool)
    if p_dropout > 0:
        keep = keep_mask(seed, offset, p_dropout, bh, sq, sk)
    rp = 1.0 / (1.0 - p_dropout) if p_dropout < 1.0 else 0.0
    pd = np.where(keep, p * (rp if p_dropout > 0 else 1.0), 0.0)
    return pd @ v, lse, p, keep


def attention_bwd(q, k, v, do, scale, causal=False, p_dropout=0.0, seed=0, offset=0, bh=0, bias=None):
    """Analytic gradients for one (b, h) in float64 (uses its own forward)."""
    o, lse, p, keep = attention_fwd(q, k, v, scale, causal, p_dropout, seed, offset, bh, bias)
    rp = (1.0 / (1.0 - p_dropout) if p_dropout < 1.0 else 0.0) if p_dropout > 0 else 1.0
    pd = np.where(keep, p * rp, 0.0)
    dv = pd.T @ do
    dp = np.where(keep, (do @ v.T) * rp, 0.0)
    D = (do * o).sum(axis=1, keepdims=True)
    ds = p * (dp - D)
    dq = scale * ds @ k
    dk = scale * ds.T @ q
    return dq, dk, dv
