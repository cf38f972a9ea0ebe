"""GPU parity of the tcgen05 forward kernel against the CPU oracle, through the C ABI.
Tolerances (BASELINE.json north_star): 2e-2 relative on out, 1e-3 on LSE."""
import numpy as np
import pytest

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL_REL = 2e-2
TOL_LSE = 1e-3

CASES = {
    # BASELINE.json configs[0]
    "c1_b1h4s128d64": lambda: cases.dense(1, 128, 128, 4, 4, 64),
    "d128_s128": lambda: cases.dense(1, 128, 128, 2, 2, 128),
    "d128_s256_causal": lambda: cases.dense(2, 256, 256, 2, 2, 128, is_causal=True),
    "d128_s384_causal": lambda: cases.dense(1, 384, 384, 3, 3, 128, is_causal=True),
    "d64_s640": lambda: cases.dense(1, 640, 640, 2, 2, 64),
    "d64_s640_causal": lambda: cases.dense(1, 640, 640, 2, 2, 64, is_causal=True),
    "ragged_300x450": lambda: cases.dense(2, 300, 450, 2, 2, 128),
    "ragged_300x450_causal": lambda: cases.dense(2, 300, 450, 2, 2, 128, is_causal=True),
    "sq_gt_sk_causal_empty_rows": lambda: cases.dense(1, 450, 300, 2, 2, 64, is_causal=True),
    "tiny_1x1": lambda: cases.dense(2, 1, 1, 2, 2, 64),
    "tiny_5x700_causal": lambda: cases.dense(1, 5, 700, 2, 1, 128, is_causal=True),
    "gqa_8to2": lambda: cases.dense(2, 257, 257, 8, 2, 128, is_causal=True),
    "fp16_d128": lambda: cases.dense(1, 512, 512, 2, 2, 128, is_causal=True, is_bf16=False),
    "fp16_d64": lambda: cases.dense(1, 200, 333, 2, 2, 64, is_bf16=False),
    "dropout_d128": lambda: cases.dense(2, 384, 384, 2, 2, 128, is_causal=True, p_dropout=0.1, seed=1234, offset=0),
    "dropout_d64_p50": lambda: cases.dense(1, 300, 450, 3, 3, 64, p_dropout=0.5, seed=2**40 + 3, offset=2**33 + 9),
    "dropout_p03_fp16": lambda: cases.dense(1, 256, 256, 2, 2, 128, p_dropout=0.03, seed=7, offset=1, is_bf16=False),
    "dropout_p80": lambda: cases.dense(1, 256, 256, 2, 2, 64, p_dropout=0.8, seed=7, offset=1),
    # head sizes between the two compiled widths ride the 64 / 128 kernels (TMA zero fill)
    "d96_causal_gqa": lambda: cases.dense(2, 300, 300, 4, 2, 96, is_causal=True),
    "d80_ragged": lambda: cases.dense(1, 200, 333, 2, 2, 80),
    "d40_dropout": lambda: cases.dense(1, 257, 257, 2, 2, 40, is_causal=True, p_dropout=0.2, seed=11, offset=4),
    "d32_fp16": lambda: cases.dense(1, 128, 384, 2, 2, 32, is_bf16=False),
    "d104_dropout": lambda: cases.dense(1, 384, 384, 2, 1, 104, p_dropout=0.1, seed=5, offset=0),
    "d8_tiny": lambda: cases.dense(1, 130, 130, 1, 1, 8, is_causal=True),
    # head sizes 136..256 ride the one-tile D = 256 kernel
    "d256_causal": lambda: cases.dense(2, 384, 384, 2, 2, 256, is_causal=True),
    "d256_ragged_gqa": lambda: cases.dense(1, 300, 450, 4, 2, 256),
    "d192_dropout": lambda: cases.dense(1, 257, 257, 2, 2, 192, is_causal=True, p_dropout=0.1, seed=3, offset=9),
    "d160_fp16": lambda: cases.dense(1, 200, 333, 2, 1, 160, is_bf16=False),
    "d136_sq_gt_sk_causal": lambda: cases.dense(1, 450, 300, 2, 2, 136, is_causal=True),
    "d224_tiny_5x700": lambda: cases.dense(1, 5, 700, 2, 1, 224, is_causal=True),
    "d256_packed": lambda: cases.varlen([200, 77, 300], [200, 77, 300], 4, 2, 256, is_causal=True),
    "d200_packed_sq_ne_sk": lambda: cases.varlen([100, 300], [260, 128], 2, 2, 200),
    "c2_small_b2h4s2048d128": lambda: cases.dense(2, 2048, 2048, 4, 4, 128, is_causal=True),
    # persistent CTAs: 512 / 640 work items on 148 SMs (both item-queue slots reused, O tiles handed from one
    # item's epilogue to the next item's first P V), with the staged (no dropout) and direct (dropout) epilogues
    "many_items_b16h8s1024d64": lambda: cases.dense(16, 1024, 1024, 8, 8, 64, is_causal=True),
    "many_items_dropout_d128": lambda: cases.dense(10, 700, 700, 16, 4, 128, is_causal=True, p_dropout=0.1, seed=21),
    "many_items_ragged_d72": lambda: cases.dense(40, 300, 450, 4, 2, 72),
    # decode-like: one query row, GQA -> the group's query heads ride one tile as rows (strided Q map)
    "decode_gqa_8to2_d128": lambda: cases.dense(3, 1, 700, 8, 2, 128),
    "decode_gqa_32to8_d128_causal": lambda: cases.dense(5, 1, 1024, 32, 8, 128, is_causal=True),
    "decode_gqa_12to4_d64_fp16": lambda: cases.dense(2, 1, 333, 12, 4, 64, is_bf16=False),
    "decode_mqa_16to1_d96": lambda: cases.dense(2, 1, 450, 16, 1, 96),
    "decode_mha_no_pack": lambda: cases.dense(2, 1, 450, 4, 4, 128),
}


def run_and_check(label, pb, seed=42, std=1.0):
    import gpu_util as G
    q, k, v, _ = cases.make_inputs(pb, seed=seed, with_dout=False, std=std)
    ref, lp_out = G.oracle_fwd(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2, "tcgen05 path did not run"
    G.check_fwd(label, pb, got, ref, lp_out)


@pytest.mark.parametrize("name", list(CASES))
def test_tc_fwd(name):
    run_and_check("tc_fwd/" + name, CASES[name]())


def test_tc_fwd_growing_max_forces_rescale():
    """Scores whose running max keeps growing across KV tiles exercise the lazy O rescale."""
    import gpu_util as G
    pb = cases.dense(1, 256, 1024, 2, 2, 128)
    q, k, v, _ = cases.make_inputs(pb, seed=3, with_dout=False)
    kf = O.from_storage(k, True)
    ramp = (1.0 + 3.0 * np.arange(1024, dtype=np.float32) / 1024.0)[:, None, None]
    k = O.to_storage(kf * ramp, True)
    qf = O.from_storage(q, True)
    q = O.to_storage(qf * 3.0, True)
    ref = O.forward(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2
    assert cases.rel_err(G.f32(got["out"], pb), ref["out_f32"]) < TOL_REL
    assert np.abs(got["lse"] - ref["lse"]).max() < 2e-3 * max(1.0, np.abs(ref["lse"]).max() / 10)


def test_tc_fwd_matches_generic_dropout_pattern():
    """Same seed/offset: the tcgen05 kernel and the all-shapes kernel drop the same entries
    (outputs agree far more tightly than the keep-rate noise would allow otherwise)."""
    import gpu_util as G
    from flash_attention_hip_b200 import api
    pb = cases.dense(1, 256, 384, 2, 2, 64, p_dropout=0.3, seed=99, offset=5)
    q, k, v, _ = cases.make_inputs(pb, seed=4, with_dout=False)
    a = G.run_fwd(pb, q, k, v)
    assert a["path"] == 2
    api.set_path(api.PATH_GENERIC)
    try:
        b = G.run_fwd(pb, q, k, v)
    finally:
        api.set_path(api.PATH_AUTO)
    assert b["path"] == 1
    assert cases.rel_err(G.f32(a["out"], pb), G.f32(b["out"], pb)) < 1.5e-2


@pytest.mark.parametrize("kind", ["dense_bias_drop", "packed", "return_softmax"])
def test_tc_fwd_negative_softmax_scale(kind):
    """softmax_scale < 0 (attention to the LEAST similar keys) on the tcgen05 forward: the scores are negated by the MMA
    (negate-A bit) and the scale made positive, so the running maximum stays a maximum.  With a bias (added after
    scaling, unaffected), dropout, packed inputs, and through the S_dmask pass."""
    import gpu_util as G
    from flash_attention_hip_b200 import api
    if kind == "dense_bias_drop":
        pb = cases.dense(2, 260, 400, 4, 2, 128, is_causal=True, p_dropout=0.2, seed=3, offset=2**33, softmax_scale=-0.09)
        rng = np.random.default_rng(8)
        pb.attn_mask = O.to_storage(rng.standard_normal((1, 4, 260, 400), dtype=np.float32), True)
    elif kind == "packed":
        pb = cases.varlen([300, 1, 129, 64], [300, 77, 257, 64], 4, 4, 64, is_causal=True, softmax_scale=-0.125, is_bf16=False)
    else:
        pb = cases.dense(1, 200, 333, 2, 1, 72, p_dropout=0.3, seed=9, softmax_scale=-0.2)
    q, k, v, _ = cases.make_inputs(pb, seed=77, with_dout=False)
    if kind == "return_softmax":
        from test_gpu_tc_return_softmax import check_plane
        check_plane("neg_scale", pb, q, k, v)
        assert api.last_path() == api.PATH_TCGEN05
        return
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == api.PATH_TCGEN05
    G.check_fwd(f"tc_fwd/negative_scale_{kind}", pb, got, O.forward(pb, q, k, v), None)
