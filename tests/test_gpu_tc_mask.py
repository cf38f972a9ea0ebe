"""attn_mask / mask_dims (reference flash_attn.h:35-36, :98-99) on the tcgen05 kernels: additive bias
in the input dtype, dims {1|b, 1|h, sq, sk}, against the CPU oracle through the C ABI.
Tolerances (north_star): 2e-2 relative on out/dq/dk/dv, 1e-3 on LSE."""
import numpy as np
import pytest

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL_REL = 2e-2


def _mask(pb, mb, mh, seed, pad_from=None, neg=-np.inf):
    rng = np.random.default_rng(seed)
    m = rng.standard_normal((mb, mh, pb.seqlen_q, pb.seqlen_k), dtype=np.float32) * 2.0
    if pad_from is not None:                     # a padding mask: keys past pad_from are invisible
        m[..., pad_from:] = neg
    return O.to_storage(m, pb.is_bf16)


CASES = {
    # name: (problem, mask batch dim, mask head dim, padded-from column or None, pad value)
    "bh_d128": (lambda: cases.dense(2, 256, 256, 2, 2, 128), 2, 2, None, 0),
    "b1_h1_broadcast_causal": (lambda: cases.dense(2, 300, 300, 4, 2, 128, is_causal=True), 1, 1, None, 0),
    "b_h1_d64_ragged": (lambda: cases.dense(2, 200, 328, 2, 2, 64), 2, 1, None, 0),
    "padding_minus_inf": (lambda: cases.dense(2, 257, 384, 2, 2, 128), 2, 1, 300, -np.inf),
    "padding_minus_1e4_fp16": (lambda: cases.dense(1, 128, 256, 2, 2, 64, is_bf16=False), 1, 1, 200, -1e4),
    "sk_not_multiple_of_8": (lambda: cases.dense(1, 130, 203, 2, 1, 128, is_causal=True), 1, 2, None, 0),
    "dropout_and_mask": (lambda: cases.dense(1, 384, 384, 2, 2, 128, p_dropout=0.2, seed=9, offset=2), 1, 2, None, 0),
    "d96_mask": (lambda: cases.dense(1, 192, 192, 2, 2, 96, is_causal=True), 1, 1, None, 0),
    # several KV tiles and items per CTA: the bias tile borrows the epilogue's staging block (forward), prefetch in the backward
    "many_tiles_bh": (lambda: cases.dense(3, 700, 1024, 4, 2, 128, is_causal=True), 3, 4, None, 0),
    "many_items_b1h1": (lambda: cases.dense(24, 520, 520, 8, 8, 64), 1, 1, None, 0),
}


@pytest.mark.parametrize("name", list(CASES))
def test_tc_attn_mask_fwd_bwd(name):
    import gpu_util as G
    mk, mb, mh, pad, neg = CASES[name]
    pb = mk()
    pb.attn_mask = _mask(pb, mb, mh, seed=21, pad_from=pad, neg=neg)
    q, k, v, dout = cases.make_inputs(pb, seed=13)
    ref = O.forward(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2, "tcgen05 forward did not run"
    assert cases.rel_err(G.f32(got["out"], pb), ref["out_f32"]) < TOL_REL
    fin = np.isfinite(ref["lse"])
    assert np.array_equal(np.isfinite(got["lse"]), fin)
    assert np.abs(got["lse"][fin] - ref["lse"][fin]).max() < 1e-3
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    assert gb["path"] == 2, "tcgen05 backward did not run"
    for t in ("dq", "dk", "dv"):
        assert cases.rel_err(G.f32(gb[t], pb), rb[t + "_f32"]) < TOL_REL, t


def test_tc_mask_fully_masked_rows_give_zero_and_inf_lse():
    import gpu_util as G
    pb = cases.dense(1, 128, 128, 2, 2, 128)
    m = np.zeros((1, 1, 128, 128), np.float32)
    m[:, :, 5:9, :] = -np.inf                    # four query rows see nothing
    pb.attn_mask = O.to_storage(m, True)
    q, k, v, _ = cases.make_inputs(pb, seed=3, with_dout=False)
    got = G.run_fwd(pb, q, k, v)
    out = G.f32(got["out"], pb).reshape(128, 2, 128)
    assert got["path"] == 2
    assert np.all(out[5:9] == 0) and np.all(np.isposinf(got["lse"][0, :, 5:9]))
    assert np.all(np.isfinite(got["lse"][0, :, :5]))
