"""GPU parity of the tcgen05 backward kernel against the CPU oracle, through the C ABI.
Tolerance (BASELINE.json north_star): 2e-2 relative on dq/dk/dv."""
import numpy as np
import pytest

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL_REL = 2e-2

CASES = {
    "d128_s128": lambda: cases.dense(1, 128, 128, 2, 2, 128),
    "d64_s128": lambda: cases.dense(1, 128, 128, 4, 4, 64),
    "d128_s256_causal": lambda: cases.dense(2, 256, 256, 2, 2, 128, is_causal=True),
    "d128_s384_causal": lambda: cases.dense(1, 384, 384, 3, 3, 128, is_causal=True),
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
    "dropout_d64_p50_gqa": lambda: cases.dense(1, 300, 450, 4, 2, 64, p_dropout=0.5, seed=2**40 + 3, offset=2**33 + 9),
    "dropout_p80": lambda: cases.dense(1, 256, 256, 2, 2, 64, p_dropout=0.8, seed=7, offset=1),
    # head sizes between the two compiled widths ride the 64 / 128 kernels (TMA zero fill)
    "d96_causal_gqa": lambda: cases.dense(2, 300, 300, 4, 2, 96, is_causal=True),
    "d80_ragged": lambda: cases.dense(1, 200, 333, 2, 2, 80),
    "d40_dropout": lambda: cases.dense(1, 257, 257, 2, 2, 40, is_causal=True, p_dropout=0.2, seed=11, offset=4),
    "d32_fp16": lambda: cases.dense(1, 128, 384, 2, 2, 32, is_bf16=False),
    "d104_dropout": lambda: cases.dense(1, 384, 384, 2, 1, 104, p_dropout=0.1, seed=5, offset=0),
    "d8_tiny": lambda: cases.dense(1, 130, 130, 1, 1, 8, is_causal=True),
    "c2_small_b2h4s2048d128": lambda: cases.dense(2, 2048, 2048, 4, 4, 128, is_causal=True, p_dropout=0.1, seed=5),
    # persistent CTAs: 640 key-tile items on 148 SMs (4-5 items per CTA: both item-queue slots reused, ring
    # phases carried across items of different lengths), and a GQA mix with 9 key tiles per sequence
    "many_items_b10h8s1024d64": lambda: cases.dense(10, 1024, 1024, 8, 8, 64, is_causal=True, p_dropout=0.1, seed=3),
    "many_items_gqa_s1152d128": lambda: cases.dense(6, 1152, 1152, 16, 4, 128, is_causal=True, p_dropout=0.2, seed=8, offset=5),
    "many_items_ragged_full": lambda: cases.dense(40, 200, 333, 4, 4, 128),
    # non-causal shapes for the ordered-dQ walk (cyclic start tile): more key tiles than query tiles and the reverse
    "det_full_s1024_gqa": lambda: cases.dense(3, 1024, 1024, 8, 2, 128, p_dropout=0.1, seed=4, offset=2),
    "det_full_sq_lt_sk": lambda: cases.dense(2, 300, 1100, 4, 4, 64),
    "det_full_sq_gt_sk": lambda: cases.dense(2, 1100, 300, 4, 2, 128),
    # head sizes 136..256: fa_tc_bwd_wide.cu (items = key tile x half of the head dimension, 64-query steps)
    "d256_s128": lambda: cases.dense(1, 128, 128, 2, 2, 256),
    "d256_s64_one_step": lambda: cases.dense(1, 64, 128, 1, 1, 256),
    "d256_causal": lambda: cases.dense(2, 384, 384, 2, 2, 256, is_causal=True),
    "d256_ragged_gqa": lambda: cases.dense(1, 300, 450, 4, 2, 256),
    "d256_causal_gqa_drop": lambda: cases.dense(2, 200, 200, 4, 2, 256, is_causal=True, p_dropout=0.1, seed=8, offset=3),
    "d192_dropout": lambda: cases.dense(1, 257, 257, 2, 2, 192, is_causal=True, p_dropout=0.1, seed=3, offset=9),
    "d192_ragged": lambda: cases.dense(1, 130, 333, 2, 2, 192),
    "d160_fp16": lambda: cases.dense(1, 200, 333, 2, 1, 160, is_bf16=False),
    "d160_sq_gt_sk_causal_fp16": lambda: cases.dense(1, 300, 150, 2, 1, 160, is_causal=True, is_bf16=False),
    "d136_sq_gt_sk_causal": lambda: cases.dense(1, 450, 300, 2, 2, 136, is_causal=True),
    "d224_tiny_5x700": lambda: cases.dense(1, 5, 700, 2, 1, 224, is_causal=True),
    "d256_packed": lambda: cases.varlen([200, 77, 300], [200, 77, 300], 4, 2, 256, is_causal=True),
    "d200_packed_sq_ne_sk": lambda: cases.varlen([100, 300], [260, 128], 2, 2, 200),
    "d136_packed": lambda: cases.varlen([100, 1, 77], [130, 64, 77], 2, 2, 136, is_causal=True),
    "d256_many_items": lambda: cases.dense(6, 640, 640, 8, 4, 256, is_causal=True, p_dropout=0.2, seed=8, offset=5),
    "d192_many_items_packed": lambda: cases.varlen([520] * 24, [520] * 24, 4, 4, 192),
}


@pytest.mark.parametrize("name", list(CASES))
def test_tc_bwd(name):
    import gpu_util as G
    pb = CASES[name]()
    q, k, v, dout = cases.make_inputs(pb, seed=42)
    ref = O.forward(pb, q, k, v)
    rb, lp = G.oracle_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    assert gb["path"] == 2, "tcgen05 backward did not run"
    G.check_bwd("tc_bwd/" + name, pb, gb, rb, lp)


def test_tc_fwd_bwd_chain():
    """Backward fed by the CUDA forward's own out/lse (what a framework does)."""
    import gpu_util as G
    pb = cases.dense(2, 512, 512, 4, 2, 128, is_causal=True, p_dropout=0.1, seed=77, offset=3)
    q, k, v, dout = cases.make_inputs(pb, seed=9)
    f = G.run_fwd(pb, q, k, v)
    gb = G.run_bwd(pb, dout, q, k, v, f["out"], f["lse"])
    rb = O.backward(pb, dout, q, k, v, f["out"], f["lse"])     # the oracle's backward of the SAME (out, lse)
    assert f["path"] == 2 and gb["path"] == 2
    G.check_bwd("tc_fwd_bwd_chain", pb, gb, rb, None)


@pytest.mark.parametrize("name", ["d128_s384_causal", "gqa_8to2", "ragged_300x450", "ragged_300x450_causal", "dropout_d128",
                                  "d96_causal_gqa", "c2_small_b2h4s2048d128", "sq_gt_sk_causal_empty_rows",
                                  "tiny_5x700_causal", "fp16_d64", "many_items_b10h8s1024d64",
                                  "many_items_gqa_s1152d128", "many_items_ragged_full", "det_full_s1024_gqa",
                                  "det_full_sq_lt_sk", "det_full_sq_gt_sk",
                                  # head sizes above 128: the two-launch pair of fa_tc_bwd_wide.cu
                                  "d256_s128", "d256_causal", "d256_ragged_gqa", "d256_causal_gqa_drop", "d192_dropout",
                                  "d160_sq_gt_sk_causal_fp16", "d136_sq_gt_sk_causal", "d224_tiny_5x700", "d256_packed",
                                  "d200_packed_sq_ne_sk", "d256_many_items", "d192_many_items_packed"])
def test_tc_bwd_deterministic_is_bit_reproducible_and_correct(name):
    """num_splits == 1 (reference flash_attn.h:94): key tiles add into dQ in a fixed order, so two runs
    agree bit for bit; the result stays within the parity tolerance."""
    import gpu_util as G
    pb = CASES[name]()
    q, k, v, dout = cases.make_inputs(pb, seed=42)
    ref = O.forward(pb, q, k, v)
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    a = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
    b = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
    assert a["path"] == 2 and b["path"] == 2, "tcgen05 backward did not run"
    for t in ("dq", "dk", "dv"):
        assert np.array_equal(a[t], b[t]), t
    G.check_bwd("tc_bwd_deterministic/" + name, pb, a, rb, None)


@pytest.mark.parametrize("variant", ["null_dq_accum", "unrounded_sizes", "d192_null_dq_accum", "packed_null_dq_accum"])
def test_tc_bwd_without_a_usable_dq_accum(variant):
    """A caller whose backward scratch is smaller than include/flash_attn.h sizes it -- dq_accum NULL, or the *_rounded
    arguments not rounded (softmax_d then has exactly seqlen_q entries per row) -- still gets the tcgen05 kernels: the
    two-launch backward needs no dq_accum (the packed dq_accum size is the one thing that
    cannot be verified against Paddle's call site here)."""
    import ctypes
    import torch
    import gpu_util as G
    from flash_attention_hip_b200 import api
    if variant == "packed_null_dq_accum":
        pb = cases.varlen([300, 77, 515], [300, 77, 515], 4, 2, 128, is_causal=True, p_dropout=0.1, seed=4, offset=1)
    else:
        pb = cases.dense(2, 300, 450, 4, 2, 192 if variant.startswith("d192") else 128, is_causal=True, p_dropout=0.1, seed=4, offset=1)
    q, k, v, dout = cases.make_inputs(pb, seed=42)
    ref = O.forward(pb, q, k, v)
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    t = [G.to_dev(a, pb.is_bf16) for a in (dout, q, k, v, ref["out"])]
    lse = torch.from_numpy(np.ascontiguousarray(ref["lse"], dtype=np.float32)).cuda()
    nb = pb.batch
    sq_r = pb.seqlen_q if variant == "unrounded_sizes" else api.round_up(pb.seqlen_q, 128)
    d_r = pb.head_size if variant == "unrounded_sizes" else api.round_up(pb.head_size, 32)
    g = G.Guarded()
    softmax_d = g.make((nb, pb.num_heads, sq_r), torch.float32)
    dq_accum = g.make((nb, sq_r, pb.num_heads, d_r), torch.float32) if variant == "unrounded_sizes" else None
    grads = [g.make(tuple(x.shape), x.dtype) for x in (t[1], t[2], t[3])]
    P = lambda x: None if x is None else ctypes.c_void_p(x.data_ptr())
    L = api.lib()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    scale = pb.softmax_scale if pb.softmax_scale else pb.head_size ** -0.5
    common = (sq_r, api.round_up(pb.seqlen_k, 128), pb.num_heads, pb.num_heads_k, pb.head_size, d_r, pb.p_dropout, scale, 1.0 / scale,
              pb.is_causal, pb.is_bf16, 0, st, pb.seed, pb.offset, None, None)
    if pb.varlen:
        cq, ck = torch.from_numpy(pb.cu_seqlens_q).cuda(), torch.from_numpy(pb.cu_seqlens_k).cuda()
        ok = L.flash_attn_varlen_bwd(P(t[0]), P(t[1]), P(t[2]), P(t[3]), P(t[4]), P(softmax_d), P(lse), P(cq), P(ck), None,
                                     P(grads[0]), P(grads[1]), P(grads[2]), P(dq_accum), nb, pb.seqlen_q, pb.seqlen_k, *common)
    else:
        ok = L.flash_attn_bwd(P(t[0]), P(t[1]), P(t[2]), P(t[3]), P(t[4]), P(softmax_d), P(lse), None, P(grads[0]), P(grads[1]),
                              P(grads[2]), P(dq_accum), nb, pb.seqlen_q, pb.seqlen_k, *common)
    assert ok, api.last_error()
    assert api.last_path() == api.PATH_TCGEN05
    torch.cuda.synchronize()
    g.check()
    gb = dict(dq=G.to_bits(grads[0]).reshape(pb.q_shape()), dk=G.to_bits(grads[1]).reshape(pb.k_shape()),
              dv=G.to_bits(grads[2]).reshape(pb.k_shape()))
    G.check_bwd("tc_bwd_short_scratch/" + variant, pb, gb, rb, None)


@pytest.mark.parametrize("scale", [-0.11, 0.0])
@pytest.mark.parametrize("d", [64, 128, 256])
def test_tc_bwd_non_positive_softmax_scale(scale, d):
    """softmax_scale <= 0 is legal arithmetic.  The backward recomputes P from the stored LSE and runs on the tcgen05
    kernels; the forward tracks its running maximum on raw scores, so a NEGATIVE scale negates them instead (the negate-A
    bit of the Q K^T instruction descriptor) and scale == 0 runs with 1e-30 (every exponential exactly 1)."""
    import gpu_util as G
    pb = cases.dense(2, 300, 333, 4, 2, d, is_causal=True, p_dropout=0.1, seed=6, offset=2, softmax_scale=scale)
    q, k, v, dout = cases.make_inputs(pb, seed=42)
    ref = O.forward(pb, q, k, v)
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    assert gb["path"] == 2
    G.check_bwd(f"tc_bwd/scale_{scale}_d{d}", pb, gb, rb, None)
    f = G.run_fwd(pb, q, k, v)
    assert f["path"] == 2
    G.check_fwd(f"fwd/scale_{scale}_d{d}", pb, f, ref, None)
