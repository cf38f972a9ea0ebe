"""Randomised parity sweep (fixed seeds): every argument the ABI has, drawn at random -- head size (8..256), lengths
(1..900, ragged), GQA ratio, dtype, causal, dropout, attn_mask (with broadcast dims), num_splits == 1, dense or packed --
forward and backward against the CPU oracle (north_star tolerances).  The hand-written cases of the other files target
known corners; this one looks for the unknown ones across the kernel families the dispatch can pick (fused, two-launch,
wide-head, all-shapes)."""
import numpy as np
import pytest

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu

N_CASES = 160


def _draw(i):
    rng = np.random.default_rng(1000 + i)
    d = int(rng.choice([8, 16, 24, 32, 40, 64, 72, 96, 128, 136, 160, 192, 200, 256]))
    many = i >= 48 and i % 4 == 0          # more work items than SMs: the persistent queues, tensor maps re-bounded per item
    hk = int(rng.choice([2, 4])) if many else int(rng.choice([1, 2, 3]))
    h = hk * int(rng.choice([1, 1, 2, 4]))
    causal = bool(rng.integers(2))
    kw = dict(is_causal=causal, is_bf16=bool(rng.integers(4) > 0))
    if rng.integers(3) == 0:
        kw.update(p_dropout=float(rng.choice([0.05, 0.1, 0.3, 0.7])), seed=int(rng.integers(1 << 62)), offset=int(rng.integers(1 << 40)))
    packed = bool(rng.integers(3) == 0)
    big = 900 if d <= 128 else 500
    if packed:
        n = int(rng.integers(6, 14)) if many else int(rng.integers(1, 5))
        lq = [int(x) for x in rng.integers(1, big, n)]
        lk = lq if rng.integers(2) else [int(x) for x in rng.integers(1, big, n)]
        pb = cases.varlen(lq, lk, h, hk, d, **kw)
    else:
        b = int(rng.integers(6, 12)) if many else int(rng.integers(1, 4))
        sq = int(rng.integers(1, big))
        sk = sq if rng.integers(2) else int(rng.integers(1, big))
        pb = cases.dense(b, sq, sk, h, hk, d, **kw)
    masked = bool(rng.integers(4) == 0)
    if masked:
        nb = pb.batch
        mb, mh = (1 if rng.integers(2) else nb), (1 if rng.integers(2) else h)
        m = rng.standard_normal((mb, mh, pb.seqlen_q, pb.seqlen_k), dtype=np.float32) * 1.5
        if rng.integers(2):
            m[..., : pb.seqlen_k // 3] += np.where(rng.random((mb, mh, 1, pb.seqlen_k // 3)) < 0.3, -np.inf, 0).astype(np.float32)
        pb.attn_mask = O.to_storage(m, pb.is_bf16)
    det = bool(rng.integers(3) == 0)
    return pb, det


@pytest.mark.parametrize("i", range(N_CASES))
def test_random_problem_fwd_bwd(i):
    import gpu_util as G
    pb, det = _draw(i)
    q, k, v, dout = cases.make_inputs(pb, seed=50 + i)
    ref = O.forward(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    tag = (f"sweep/{i}:d{pb.head_size}_h{pb.num_heads}/{pb.num_heads_k}_{'packed' if pb.varlen else 'dense'}"
           f"{'_causal' if pb.is_causal else ''}{'_drop' if pb.p_dropout > 0 else ''}{'_mask' if pb.attn_mask is not None else ''}"
           f"{'_det' if det else ''}")
    G.check_fwd(tag, pb, got, ref, None)
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=det)
    G.check_bwd(tag, pb, gb, rb, None)
    if det:
        gb2 = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
        for t in ("dq", "dk", "dv"):
            assert np.array_equal(gb[t], gb2[t]), (tag, t)
