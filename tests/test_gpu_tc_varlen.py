"""GPU parity of the tcgen05 kernels on PACKED inputs (flash_attn_varlen_fwd / _bwd,
reference flash_attn.h:38-66, :101-134) against the CPU oracle, through the C ABI.
BASELINE.json configs[3]: mixed 128-8192 tokens, D=64/128, GQA heads."""
import numpy as np
import pytest

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL_REL = 2e-2
TOL_LSE = 1e-3

C4_LENS = [128, 8192, 512, 4096, 1000, 2048, 256, 6000]

CASES = {
    "two_seqs_d128": lambda: cases.varlen([128, 256], [128, 256], 2, 2, 128),
    "ragged_d64_causal": lambda: cases.varlen([5, 300, 129, 1], [5, 300, 129, 1], 4, 2, 64, is_causal=True),
    "ragged_d128_causal_gqa": lambda: cases.varlen([200, 77, 640], [200, 77, 640], 8, 2, 128, is_causal=True),
    "empty_seq_inside": lambda: cases.varlen([130, 0, 257], [130, 0, 257], 2, 2, 128, is_causal=True),
    "sq_ne_sk": lambda: cases.varlen([100, 300, 17], [260, 128, 400], 4, 4, 64),
    "sq_ne_sk_causal": lambda: cases.varlen([100, 300, 17], [260, 128, 400], 4, 4, 128, is_causal=True),
    "dropout_causal": lambda: cases.varlen([384, 100, 513], [384, 100, 513], 4, 2, 128, is_causal=True,
                                           p_dropout=0.1, seed=1234, offset=7),
    "fp16_dropout": lambda: cases.varlen([250, 250], [250, 250], 2, 1, 64, p_dropout=0.25, seed=3, offset=0,
                                         is_bf16=False),
    "d96_ragged_causal_gqa": lambda: cases.varlen([200, 77, 300], [200, 77, 300], 4, 2, 96, is_causal=True),
    "d48_sq_ne_sk": lambda: cases.varlen([100, 300], [260, 128], 2, 2, 48),
    "c4_mix_d128_causal_h8": lambda: cases.varlen(C4_LENS, C4_LENS, 8, 2, 128, is_causal=True),
    "c4_mix_d64_h4": lambda: cases.varlen(C4_LENS[2:], C4_LENS[2:], 4, 1, 64),
}


def check_lse(got, ref, pb):
    for b, _, sq, _, _ in cases.seq_slices(pb):
        g, r = got[b, :, :sq], ref[b, :, :sq]
        assert np.array_equal(np.isposinf(g), np.isposinf(r))
        fin = np.isfinite(r)
        assert np.abs(g[fin] - r[fin]).max(initial=0) < TOL_LSE


@pytest.mark.parametrize("name", list(CASES))
def test_tc_varlen_fwd_bwd(name):
    import gpu_util as G
    pb = CASES[name]()
    q, k, v, dout = cases.make_inputs(pb, seed=11)
    big = pb.total_q > 8000            # the C4 mixes: one oracle pass is enough CPU time
    if big:
        ref, lp_out = O.forward(pb, q, k, v), None
    else:
        ref, lp_out = G.oracle_fwd(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v)
    assert got["path"] == 2, "tcgen05 forward did not run"
    G.check_fwd("tc_varlen/" + name, pb, got, ref, lp_out)
    if big:
        rb, lp = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"]), None
    else:
        rb, lp = G.oracle_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"])
    assert gb["path"] == 2, "tcgen05 backward did not run"
    G.check_bwd("tc_varlen/" + name, pb, gb, rb, lp)


def test_varlen_equals_dense_per_sequence():
    """Packing must not change a sequence's result: each packed sequence equals the same
    sequence run alone through the dense entry point (bit-exact, same kernel, same tiles)."""
    import gpu_util as G
    lens = [300, 128, 515]
    pb = cases.varlen(lens, lens, 4, 2, 128, is_causal=True)
    q, k, v, _ = cases.make_inputs(pb, seed=5, with_dout=False)
    got = G.run_fwd(pb, q, k, v)
    for b, q0, sq, k0, sk in cases.seq_slices(pb):
        one = cases.dense(1, sq, sk, 4, 2, 128, is_causal=True)
        g1 = G.run_fwd(one, q[q0:q0 + sq], k[k0:k0 + sk], v[k0:k0 + sk])
        assert np.array_equal(g1["out"], got["out"][q0:q0 + sq])
        assert np.array_equal(g1["lse"][0], got["lse"][b, :, :sq])


def test_varlen_neighbour_poison():
    """Rows of the NEXT sequence must never leak into a ragged last tile: poison them with
    inf/NaN-free huge values in V and check the short sequence is unchanged."""
    import gpu_util as G
    lens = [100, 200]
    pb = cases.varlen(lens, lens, 2, 2, 128)
    q, k, v, _ = cases.make_inputs(pb, seed=6, with_dout=False)
    a = G.run_fwd(pb, q, k, v)
    v2 = v.copy()
    v2[100:] = O.to_storage(np.full(v2[100:].shape, np.inf, np.float32), True)
    b = G.run_fwd(pb, q, k, v2)
    assert np.array_equal(a["out"][:100], b["out"][:100])


@pytest.mark.parametrize("name", ["ragged_d128_causal_gqa", "sq_ne_sk", "sq_ne_sk_causal", "dropout_causal", "empty_seq_inside",
                                  "c4_mix_d128_causal_h8"])
def test_tc_varlen_bwd_deterministic(name):
    """num_splits == 1 on packed inputs: bit-reproducible and within the parity tolerance (the ordered-dQ turn
    ranks are computed from each sequence's own lengths)."""
    import gpu_util as G
    pb = CASES[name]()
    q, k, v, dout = cases.make_inputs(pb, seed=11)
    ref = O.forward(pb, q, k, v)
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    a = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
    b = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
    assert a["path"] == 2 and b["path"] == 2
    for t in ("dq", "dk", "dv"):
        assert np.array_equal(a[t], b[t]), t
    G.check_bwd("tc_varlen_deterministic/" + name, pb, a, rb, None)


@pytest.mark.parametrize("d", [128, 256])
def test_deterministic_backward_runs_of_empty_query_tiles(d):
    """Packed causal sequences with far fewer keys than queries: most query tiles of such a sequence see no key, so the
    mirrored dQ pass of the deterministic backward meets long RUNS of items without steps in its persistent queue.  Their
    acc_done commits used to run two phases ahead of the drain warps' parity waits (a hang, once in ~10 processes, found by
    the random sweep); repeated here with the work that made it likely (attn_mask, GQA, >148 items)."""
    import gpu_util as G
    lq = [308, 898, 318, 321, 700, 876, 230, 303, 892]
    lk = [695, 745, 753, 852, 692, 17, 473, 817, 226]
    pb = cases.varlen(lq, lk, 16, 4, d, is_causal=True)
    rng = np.random.default_rng(3)
    pb.attn_mask = O.to_storage(rng.standard_normal((1, 16, pb.seqlen_q, pb.seqlen_k), dtype=np.float32), True)
    q, k, v, dout = cases.make_inputs(pb, seed=21)
    ref = O.forward(pb, q, k, v)
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    small = cases.varlen([518, 497], [497, 518], 3, 3, 16, is_causal=True)        # a few-item launch in between, as in the sweep
    qs, ks, vs, ds_ = cases.make_inputs(small, seed=22)
    rs = O.forward(small, qs, ks, vs)
    first = None
    for rep in range(12):
        G.run_bwd(small, ds_, qs, ks, vs, rs["out"], rs["lse"], deterministic=True)
        gb = G.run_bwd(pb, dout, q, k, v, ref["out"], ref["lse"], deterministic=True)
        assert gb["path"] == 2
        if first is None:
            first = gb
            G.check_bwd(f"tc_varlen_deterministic/empty_query_tile_runs_d{d}", pb, gb, rb, None)
        else:
            for t in ("dq", "dk", "dv"):
                assert np.array_equal(gb[t], first[t]), (rep, t)
