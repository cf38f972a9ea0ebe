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
    # head sizes above 128: forward on the D = 256 build, backward as the two-launch pair of fa_tc_bwd_wide.cu
    "d256_bh_causal": (lambda: cases.dense(2, 300, 300, 4, 2, 256, is_causal=True), 2, 4, None, 0),
    "d192_b1h1_ragged": (lambda: cases.dense(2, 200, 328, 2, 2, 192), 1, 1, None, 0),
    "d256_padding_minus_inf_dropout": (lambda: cases.dense(1, 257, 384, 2, 1, 256, p_dropout=0.2, seed=4, offset=6), 1, 2, 300, -np.inf),
    "d160_sk_not_multiple_of_8_fp16": (lambda: cases.dense(1, 130, 203, 2, 1, 160, is_causal=True, is_bf16=False), 1, 2, None, 0),
    "d256_many_items": (lambda: cases.dense(10, 520, 520, 8, 4, 256), 1, 1, None, 0),
}


@pytest.mark.parametrize("name", list(CASES))
def test_tc_attn_mask_fwd_bwd(name):
    import gpu_util as G
    mk, mb, mh, pad, neg = CASES[name]
    pb = mk()
    pb.attn_mask = _mask(pb, mb, mh, seed=21, pad_from=pad, neg=neg)
    q, k, v, dout = cases.make_inputs(pb, seed=13)
    ref, lp_out = G.oracle_fwd(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2, "tcgen05 forward did not run"
    G.check_fwd("tc_mask/" + name, pb, got, ref, lp_out)
    rb, lp = G.oracle_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    assert gb["path"] == 2, "tcgen05 backward did not run"
    G.check_bwd("tc_mask/" + name, pb, gb, rb, lp)


@pytest.mark.parametrize("is_bf16", [True, False])
@pytest.mark.parametrize("d,sk", [(128, 128), (64, 128), (64, 2048), (128, 2048), (64, 1600)])
def test_tc_mask_fully_masked_rows_give_zero_and_inf_lse(d, sk, is_bf16):
    """include/flash_attn.h: a row whose bias is all -inf gives out = 0 and LSE = +inf -- for every head size and
    key length (the D = 64 builds compute part of the exponentials on the FMA pipe from 1536 keys up, where a
    clamped polynomial would turn exp2(-inf) into 2^-125; masked builds must not use it)."""
    import gpu_util as G
    pb = cases.dense(1, 128, sk, 2, 2, d, is_bf16=is_bf16)
    m = np.zeros((1, 1, 128, sk), np.float32)
    m[:, :, 5:9, :] = -np.inf                    # four query rows see nothing
    m[:, :, 40:50, sk // 2:] = -np.inf           # and ten see only the first half
    pb.attn_mask = O.to_storage(m, is_bf16)
    q, k, v, _ = cases.make_inputs(pb, seed=3, with_dout=False)
    ref, lp_out = G.oracle_fwd(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    out = G.f32(got["out"], pb).reshape(128, 2, d)
    assert got["path"] == 2
    assert np.all(out[5:9] == 0) and np.all(np.isposinf(got["lse"][0, :, 5:9]))
    assert np.all(np.isfinite(got["lse"][0, :, :5]))
    G.check_fwd(f"tc_mask/fully_masked_d{d}_sk{sk}_{'bf16' if is_bf16 else 'fp16'}", pb, got, ref, lp_out)


@pytest.mark.parametrize("causal,sq,sk", [(False, 256, 2048), (True, 2048, 1664)])
def test_tc_mask_minus_inf_key_padding_d64_long(causal, sq, sk):
    """-inf key padding at D = 64 with >= 1536 keys (the FMA-pipe exp2 switch), dense causal with sq > sk too:
    some query rows end up with no visible key at all."""
    import gpu_util as G
    pb = cases.dense(2, sq, sk, 4, 2, 64, is_causal=causal)
    rng = np.random.default_rng(5)
    m = rng.standard_normal((2, 1, sq, sk), dtype=np.float32)
    m[0, :, :, sk - 300:] = -np.inf              # batch 0: the last 300 keys are padding
    m[1, :, :, :sk - 100] = -np.inf              # batch 1: only the last 100 keys are real
    pb.attn_mask = O.to_storage(m, True)
    q, k, v, dout = cases.make_inputs(pb, seed=4)
    ref, lp_out = G.oracle_fwd(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2
    G.check_fwd(f"tc_mask/minus_inf_padding_d64_{'causal' if causal else 'full'}", pb, got, ref, lp_out)
    dead = np.isposinf(ref["lse"])
    if causal:
        assert dead.any()                        # rows above the bottom-right aligned diagonal of batch 0's real keys
    outf = G.f32(got["out"], pb).reshape(2, sq, 4, 64)
    assert np.all(outf.transpose(0, 2, 1, 3)[dead] == 0)
    rb, lp = G.oracle_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    assert gb["path"] == 2
    G.check_bwd(f"tc_mask/minus_inf_padding_d64_{'causal' if causal else 'full'}", pb, gb, rb, lp)


# attn_mask on the PACKED entry points (reference flash_attn.h:65-66, :133-134).  Layout: the
# bias is padded, [1|b, 1|h, max_seqlen_q, max_seqlen_k], entry (row, col) counted from each sequence's own start.
PACKED = {
    # name: (lens_q, lens_k, h, hk, d, kwargs, mask batch dim, mask head dim, -inf padding of the key tail)
    "mix_d128_causal": ([300, 128, 517, 1], [300, 128, 517, 1], 4, 2, 128, dict(is_causal=True), 4, 1, False),
    "cross_d64_bh": ([200, 77, 640], [328, 512, 130], 2, 2, 64, dict(), 3, 2, False),
    "broadcast_b1h1_dropout": ([256, 384, 130], [256, 384, 130], 4, 4, 128, dict(p_dropout=0.15, seed=3, offset=8), 1, 1, False),
    "key_padding_minus_inf_fp16": ([1024, 700, 256], [1024, 700, 256], 4, 1, 64, dict(is_bf16=False), 3, 1, True),
    "many_items": ([520] * 40, [520] * 40, 8, 8, 64, dict(is_causal=True), 1, 8, False),
    "wide_d256_causal": ([300, 128, 417], [300, 128, 417], 4, 2, 256, dict(is_causal=True), 3, 1, False),
    "wide_d192_cross": ([200, 77, 340], [328, 212, 130], 2, 2, 192, dict(p_dropout=0.1, seed=2, offset=3), 1, 2, False),
}


@pytest.mark.parametrize("name", list(PACKED))
def test_tc_attn_mask_packed_fwd_bwd(name):
    import gpu_util as G
    lq, lk, h, hk, d, kw, mb, mh, pad = PACKED[name]
    pb = cases.varlen(lq, lk, h, hk, d, **kw)
    rng = np.random.default_rng(31)
    m = rng.standard_normal((mb, mh, pb.seqlen_q, pb.seqlen_k), dtype=np.float32) * 2.0
    if pad:                                      # the last third of every sequence's keys is invisible
        for b in range(mb):
            m[b, :, :, lk[b] - lk[b] // 3:] = -np.inf
    pb.attn_mask = O.to_storage(m, pb.is_bf16)
    q, k, v, dout = cases.make_inputs(pb, seed=17)
    ref, lp_out = G.oracle_fwd(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2, "tcgen05 forward did not run"
    G.check_fwd("tc_mask_packed/" + name, pb, got, ref, lp_out)
    rb, lp = G.oracle_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    assert gb["path"] == 2, "tcgen05 backward did not run"
    G.check_bwd("tc_mask_packed/" + name, pb, gb, rb, lp)


def test_packed_mask_equals_dense_mask_per_sequence():
    """One packed call with a padded bias == one dense call per sequence with that sequence's corner of the bias."""
    import torch
    import gpu_util as G
    from flash_attention_hip_b200 import api
    lens = [384, 200, 129]
    pb = cases.varlen(lens, lens, 4, 2, 128, is_causal=True)
    rng = np.random.default_rng(8)
    m = O.to_storage(rng.standard_normal((3, 4, 384, 384), dtype=np.float32), True)
    pb.attn_mask = m
    q, k, v, _ = cases.make_inputs(pb, seed=2, with_dout=False)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2
    packed_out = got["out"].reshape(-1, 4, 128)
    for b, q0, sq, k0, sk in cases.seq_slices(pb):
        one = cases.dense(1, sq, sk, 4, 2, 128, is_causal=True)
        one.attn_mask = np.ascontiguousarray(m[b:b + 1, :, :sq, :sk])
        g1 = G.run_fwd(one, q.reshape(-1, 4, 128)[q0:q0 + sq], k.reshape(-1, 2, 128)[k0:k0 + sk],
                       v.reshape(-1, 2, 128)[k0:k0 + sk])
        assert np.array_equal(g1["out"].reshape(sq, 4, 128), packed_out[q0:q0 + sq]), f"sequence {b}"
        assert np.array_equal(g1["lse"][0], got["lse"][b, :, :sq])


@pytest.mark.parametrize("name", ["bh_d128", "b1_h1_broadcast_causal", "b_h1_d64_ragged", "padding_minus_inf", "sk_not_multiple_of_8",
                                  "dropout_and_mask", "many_tiles_bh", "packed:mix_d128_causal", "packed:cross_d64_bh",
                                  "d256_bh_causal", "d192_b1h1_ragged", "d160_sk_not_multiple_of_8_fp16", "packed:wide_d256_causal"])
def test_tc_attn_mask_deterministic_bwd(name):
    """attn_mask together with num_splits == 1: the two-launch deterministic backward (dK/dV, then the mirrored dQ pass,
    which reads the bias row-wise) on the tcgen05 kernels: bit-reproducible and within the parity tolerance."""
    import gpu_util as G
    if name.startswith("packed:"):
        lq, lk, h, hk, d, kw, mb, mh, pad = PACKED[name[7:]]
        pb = cases.varlen(lq, lk, h, hk, d, **kw)
        rng = np.random.default_rng(31)
        pb.attn_mask = O.to_storage(rng.standard_normal((mb, mh, pb.seqlen_q, pb.seqlen_k), dtype=np.float32) * 2.0, pb.is_bf16)
    else:
        mk, mb, mh, pad, neg = CASES[name]
        pb = mk()
        pb.attn_mask = _mask(pb, mb, mh, seed=21, pad_from=pad, neg=neg)
    q, k, v, dout = cases.make_inputs(pb, seed=13)
    ref = O.forward(pb, q, k, v)
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    a = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
    b = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
    assert a["path"] == 2 and b["path"] == 2, "tcgen05 backward did not run"
    for t in ("dq", "dk", "dv"):
        assert np.array_equal(a[t], b[t]), t
    G.check_bwd("tc_mask_deterministic/" + name, pb, a, rb, None)
