"""Dropout masks of the tcgen05 kernels, compared BIT FOR BIT with the written specification
(csrc/philox_spec.h through oracle.dropout_mask) -- north_star: "Dropout masks must be bit-exact for a
given seed/offset".  The ABI's own window on the mask (`return_softmax`) is written by a separate pass
(tests/test_gpu_tc_return_softmax.py), so the kernels that ship are probed through their ordinary outputs here:

  forward    V = identity block at key tile j (zeros elsewhere)  =>  out[row, c] = P_drop[row, jD + c] / (1-p),
             so  out != 0  <=>  the entry is visible AND kept.
  backward   dO = identity block at query tile i                 =>  dV[n, c] = P_drop[iD + c, n] / (1-p),
             which shows the mask of the TRANSPOSED path (Philox drawn per query row, 16x16 bit transposes,
             PRMT masks on P^T);
             and, with V[:, 0] = 1, dO[:, 0] = 1 (dP = 1 everywhere) and K = identity block at key tile j,
             dQ[row, c] = scale * P[row, jD+c] * (keep/(1-p) - kappa_row): its SIGN is the keep bit the dP path used.

Every (seed, offset) pair has high words set; p covers thresholds on both sides of 128; dense and packed
layouts, causal or not, D = 64 and 128, batch_base != 0.
"""
import numpy as np
import pytest

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu

SEEDS = [(1234, 0), (2**40 + 3, 2**33 + 9), (0xDEADBEEFCAFEF00D, 0xFFFFFFFF00000001)]


def _visible(pb, sq, sk):
    """[sq, sk] bool: causal visibility (bottom-right aligned)."""
    if not pb.is_causal:
        return np.ones((sq, sk), bool)
    i = np.arange(sq)[:, None]
    j = np.arange(sk)[None, :]
    return j <= i + (sk - sq)


def _expected_keep(pb, b, h, sq, sk):
    bh = (pb.batch_base + b) * pb.num_heads + h
    return O.dropout_mask(pb.seed, pb.offset, pb.p_dropout, bh, sq, sk).astype(bool)


def _small_inputs(pb, seed):
    # small scores: P stays far from underflow, so "nonzero" can only mean "kept and visible"
    return cases.make_inputs(pb, seed=seed, std=0.25)


FWD_CASES = {
    "d128_p10_causal": lambda s, o: cases.dense(2, 300, 384, 4, 2, 128, is_causal=True, p_dropout=0.1, seed=s, offset=o),
    "d128_p50": lambda s, o: cases.dense(1, 257, 640, 2, 2, 128, p_dropout=0.5, seed=s, offset=o),
    "d128_p03": lambda s, o: cases.dense(1, 130, 256, 2, 1, 128, p_dropout=0.03, seed=s, offset=o),
    "d128_p80_batch_base": lambda s, o: cases.dense(2, 200, 256, 2, 2, 128, p_dropout=0.8, seed=s, offset=o, batch_base=5),
    "d64_p10_causal": lambda s, o: cases.dense(1, 333, 333, 4, 4, 64, is_causal=True, p_dropout=0.1, seed=s, offset=o),
    "d64_p50_sq_gt_sk": lambda s, o: cases.dense(1, 450, 300, 2, 2, 64, is_causal=True, p_dropout=0.5, seed=s, offset=o),
    "packed_d128_p10_causal": lambda s, o: cases.varlen([200, 77, 300], [200, 77, 300], 4, 2, 128, is_causal=True,
                                                        p_dropout=0.1, seed=s, offset=o),
    "packed_d64_p25_sq_ne_sk": lambda s, o: cases.varlen([100, 300], [260, 128], 2, 2, 64, p_dropout=0.25, seed=s, offset=o,
                                                          batch_base=3),
    # many work items per persistent CTA (the draw-ahead state is carried from item to item)
    "many_items_d128_p10": lambda s, o: cases.dense(12, 512, 512, 16, 4, 128, is_causal=True, p_dropout=0.1, seed=s, offset=o),
}


@pytest.mark.parametrize("si", range(len(SEEDS)))
@pytest.mark.parametrize("name", list(FWD_CASES))
def test_forward_mask_is_the_specification(name, si):
    import gpu_util as G
    if name.startswith("many_items") and si:
        pytest.skip("one seed is enough for the big case")
    pb = FWD_CASES[name](*SEEDS[si])
    d = pb.head_size
    q, k, _, _ = _small_inputs(pb, seed=31)
    n_tiles = (pb.seqlen_k + d - 1) // d
    checked = 0
    for j in range(n_tiles):
        # V = identity at keys [j d, j d + d) of EVERY sequence, zero elsewhere
        vf = np.zeros(pb.k_shape(), np.float32)
        for b, _, _, k0, sk in cases.seq_slices(pb):
            for c in range(min(d, sk - j * d)):
                vf[k0 + j * d + c, :, c] = 1.0
        got = G.run_fwd(pb, q, k, O.to_storage(vf, pb.is_bf16))
        assert got["path"] == 2, "tcgen05 forward did not run"
        out = G.f32(got["out"], pb)
        for b, q0, sq, k0, sk in cases.seq_slices(pb):
            w = min(d, sk - j * d)
            if w <= 0 or sq == 0:
                continue
            vis = _visible(pb, sq, sk)[:, j * d:j * d + w]
            for h in range(pb.num_heads):
                if name.startswith("many_items") and (b * pb.num_heads + h) % 7:
                    continue
                keep = _expected_keep(pb, b, h, sq, sk)[:, j * d:j * d + w]
                nz = out[q0:q0 + sq, h, :w] != 0
                assert np.array_equal(nz, keep & vis), \
                    f"{name}: batch {b} head {h} key tile {j}: {np.count_nonzero(nz != (keep & vis))} mask bits differ"
                assert not out[q0:q0 + sq, h, w:].any()
                checked += int(vis.sum())
    assert checked > 0


BWD_CASES = {
    "d128_p10_causal": lambda s, o: cases.dense(2, 384, 300, 2, 2, 128, is_causal=True, p_dropout=0.1, seed=s, offset=o),
    "d128_p50": lambda s, o: cases.dense(1, 257, 450, 2, 2, 128, p_dropout=0.5, seed=s, offset=o),
    "d128_p80_batch_base": lambda s, o: cases.dense(2, 256, 200, 2, 2, 128, p_dropout=0.8, seed=s, offset=o, batch_base=5),
    "d64_p10_causal": lambda s, o: cases.dense(1, 333, 333, 3, 3, 64, is_causal=True, p_dropout=0.1, seed=s, offset=o),
    "d64_p03": lambda s, o: cases.dense(1, 130, 200, 2, 2, 64, p_dropout=0.03, seed=s, offset=o),
    "packed_d128_p10_causal": lambda s, o: cases.varlen([200, 77, 300], [200, 77, 300], 2, 2, 128, is_causal=True,
                                                        p_dropout=0.1, seed=s, offset=o),
    "packed_d64_p25_sq_ne_sk": lambda s, o: cases.varlen([100, 300], [260, 128], 2, 2, 64, p_dropout=0.25, seed=s, offset=o,
                                                          batch_base=3),
    "many_items_d128_p10": lambda s, o: cases.dense(10, 512, 1024, 4, 4, 128, is_causal=True, p_dropout=0.1, seed=s, offset=o),
    # head sizes above 128: fa_tc_bwd_wide.cu (two Philox calls per thread and step, one-word bit transposes)
    "d256_p10_causal": lambda s, o: cases.dense(2, 384, 300, 2, 2, 256, is_causal=True, p_dropout=0.1, seed=s, offset=o),
    "d192_p50_batch_base": lambda s, o: cases.dense(1, 257, 450, 2, 2, 192, p_dropout=0.5, seed=s, offset=o, batch_base=4),
    "packed_d256_p25": lambda s, o: cases.varlen([200, 77, 300], [260, 77, 130], 2, 2, 256, p_dropout=0.25, seed=s, offset=o),
}


@pytest.mark.parametrize("si", range(len(SEEDS)))
@pytest.mark.parametrize("name", list(BWD_CASES))
def test_backward_transposed_mask_is_the_specification(name, si):
    """dV = P_drop^T dO with dO = identity rows: dV[n, c] != 0  <=>  (row i d + c, key n) visible and kept.
    num_heads == num_heads_k in these cases, so no other head adds into the same dV."""
    import gpu_util as G
    if name.startswith("many_items") and si:
        pytest.skip("one seed is enough for the big case")
    pb = BWD_CASES[name](*SEEDS[si])
    assert pb.num_heads == pb.num_heads_k
    d = pb.head_size
    q, k, v, _ = _small_inputs(pb, seed=32)
    ref = O.forward(pb, q, k, v)
    n_tiles = (pb.seqlen_q + d - 1) // d
    checked = 0
    for i in range(n_tiles):
        dof = np.zeros(pb.q_shape(), np.float32)
        for b, q0, sq, _, _ in cases.seq_slices(pb):
            for c in range(min(d, sq - i * d)):
                dof[q0 + i * d + c, :, c] = 1.0
        gb = G.run_bwd(pb, O.to_storage(dof, pb.is_bf16), q, k, v, ref["out"], ref["lse"])
        assert gb["path"] == 2, "tcgen05 backward did not run"
        dv = G.f32(gb["dv"], pb)
        for b, q0, sq, k0, sk in cases.seq_slices(pb):
            w = min(d, sq - i * d)
            if w <= 0 or sk == 0:
                continue
            vis = _visible(pb, sq, sk)[i * d:i * d + w]              # [w rows, sk]
            for h in range(pb.num_heads):
                if name.startswith("many_items") and (b * pb.num_heads + h) % 5:
                    continue
                keep = _expected_keep(pb, b, h, sq, sk)[i * d:i * d + w]
                nz = dv[k0:k0 + sk, h, :w] != 0                       # [sk, w]
                assert np.array_equal(nz, (keep & vis).T), \
                    f"{name}: batch {b} head {h} query tile {i}: {np.count_nonzero(nz != (keep & vis).T)} mask bits differ"
                assert not dv[k0:k0 + sk, h, w:].any()
                checked += int(vis.sum())
    assert checked > 0


@pytest.mark.parametrize("si", range(len(SEEDS)))
@pytest.mark.parametrize("name", ["d128_p10", "d128_p50", "d64_p80", "packed_d128_p25", "d256_p25"])
def test_backward_dp_mask_sign_probe(name, si):
    """The dP path of the backward applies the same keep bits with test + select.  With dP == 1 everywhere
    (V[:, 0] = 1, dO[:, 0] = 1) and K = identity block at key tile j:
        dQ[row, c] = scale * P[row, j d + c] * (keep / (1-p) - kappa_row),  kappa_row = out[row, 0] in (0, 1/(1-p)),
    so a kept entry gives dQ > 0 and a dropped one dQ < 0.  Rows whose kappa leaves no margin are skipped."""
    import gpu_util as G
    s, o = SEEDS[si]
    pb = {"d128_p10": lambda: cases.dense(1, 300, 384, 2, 2, 128, p_dropout=0.1, seed=s, offset=o),
          "d128_p50": lambda: cases.dense(2, 257, 256, 2, 2, 128, p_dropout=0.5, seed=s, offset=o, batch_base=2),
          "d64_p80": lambda: cases.dense(1, 200, 320, 2, 2, 64, p_dropout=0.8, seed=s, offset=o),
          "d256_p25": lambda: cases.dense(1, 200, 512, 2, 2, 256, p_dropout=0.25, seed=s, offset=o),
          "packed_d128_p25": lambda: cases.varlen([150, 260], [256, 384], 2, 2, 128, p_dropout=0.25, seed=s, offset=o)}[name]()
    d = pb.head_size
    rp = 1.0 / (1.0 - pb.p_dropout)
    rng = np.random.default_rng(33)
    q = O.to_storage(rng.standard_normal(pb.q_shape(), dtype=np.float32) * 0.25, pb.is_bf16)
    vf = np.zeros(pb.k_shape(), np.float32); vf[:, :, 0] = 1.0
    dof = np.zeros(pb.q_shape(), np.float32); dof[:, :, 0] = 1.0
    v, dout = O.to_storage(vf, pb.is_bf16), O.to_storage(dof, pb.is_bf16)
    checked = 0
    for j in range((pb.seqlen_k + d - 1) // d):
        kf = np.zeros(pb.k_shape(), np.float32)
        for b, _, _, k0, sk in cases.seq_slices(pb):
            for c in range(min(d, sk - j * d)):
                kf[k0 + j * d + c, :, c] = 1.0
        k = O.to_storage(kf, pb.is_bf16)
        ref = O.forward(pb, q, k, v)
        gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
        assert gb["path"] == 2, "tcgen05 backward did not run"
        dq = G.f32(gb["dq"], pb)
        kappa = O.from_storage(ref["out"], pb.is_bf16)[:, :, 0]          # what the kernel's D = rowsum(dO o O) sees
        for b, q0, sq, k0, sk in cases.seq_slices(pb):
            w = min(d, sk - j * d)
            if w <= 0:
                continue
            for h in range(pb.num_heads):
                keep = _expected_keep(pb, b, h, sq, sk)[:, j * d:j * d + w]
                kap = kappa[q0:q0 + sq, h]
                rows = (kap > 0.05) & (kap < rp - 0.05)                  # both signs have a margin
                got = dq[q0:q0 + sq, h, :w][rows]
                assert np.array_equal(got > 0, keep[rows]) and np.array_equal(got < 0, ~keep[rows]), \
                    f"{name}: batch {b} head {h} key tile {j}: dP-path mask bits differ"
                checked += int(rows.sum()) * w
    assert checked > 1000
