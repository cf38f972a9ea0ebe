"""CPU-only: pins the oracle (oracle/attn_oracle.c).

The reference has no fixtures for this path, so the
oracle is pinned by (1) the published Random123 Philox4x32 known-answer
vectors for 7 rounds (the dropout specification, csrc/philox_spec.h) and 10 rounds, (2) agreement of its fp32 build with its fp64 build and with an
independent numpy float64 restatement, (3) finite differences, and
(4) committed golden outputs under tests/golden/.
"""
import os

import numpy as np
import pytest

import cases
import ref_numpy as R
from oracle import oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

# Random123 kat_vectors, philox4x32 with 7 and 10 rounds: (rounds, counter, key, expected)
_PI_CTR, _PI_KEY = [0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0]
PHILOX_KAT = [
    (7, [0, 0, 0, 0], [0, 0], [0x5F6FB709, 0x0D893F64, 0x4F121F81, 0x4F730A48]),
    (7, [0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x5207DDC2, 0x45165E59, 0x4D8EE751, 0x8C52F662]),
    (7, _PI_CTR, _PI_KEY, [0x4DFCCABA, 0x190A87F0, 0xC47362BA, 0xB6B5242A]),
    (10, [0, 0, 0, 0], [0, 0], [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]),
    (10, [0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2, [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]),
    (10, _PI_CTR, _PI_KEY, [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]),
]


@pytest.mark.parametrize("rounds,ctr,key,exp", PHILOX_KAT)
def test_philox_known_answers(rounds, ctr, key, exp):
    assert list(O.philox(ctr, key, rounds)) == exp
    assert R.philox4x32(ctr, key, rounds) == exp
    if rounds == O.philox_rounds():          # the generator the dropout specification uses
        assert list(O.philox(ctr, key)) == exp


def test_dropout_spec_round_count():
    assert O.philox_rounds() == 7 == R.ROUNDS


def test_storage_roundtrip_matches_numpy_half():
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.standard_normal(4000).astype(np.float32) * s
                        for s in (1e-8, 1e-5, 1e-3, 1.0, 100.0, 7e4)])
    x = np.concatenate([x, np.array([0.0, -0.0, 65504.0, 65520.0, 2.0 ** -24, 2.0 ** -25,
                                     np.float32(2.0 ** -25) * np.float32(1.0000001)], np.float32)])
    lib = O.lib()
    got = np.array([lib.fa_oracle_f32_to_storage(float(v), 0) for v in x], np.uint16)
    assert np.array_equal(got, x.astype(np.float16).view(np.uint16))
    gotb = np.array([lib.fa_oracle_f32_to_storage(float(v), 1) for v in x], np.uint16)
    assert np.array_equal(gotb, O.to_storage(x, True))
    back = np.array([lib.fa_oracle_storage_to_f32(int(v), 0) for v in got], np.float32)
    assert np.array_equal(back, got.view(np.float16).astype(np.float32))


def test_dropout_mask_matches_numpy_and_rate():
    for (seed, offset, p, bh, sq, sk) in [(1234, 0, 0.1, 0, 64, 200), (2**40 + 7, 2**33 + 5, 0.5, 17, 33, 47),
                                          (3, 99, 0.9, 5, 16, 16)]:
        a = O.dropout_mask(seed, offset, p, bh, sq, sk).astype(bool)
        b = R.keep_mask(seed, offset, p, bh, sq, sk)
        assert np.array_equal(a, b)
    big = O.dropout_mask(1234, 0, 0.1, 3, 512, 512)
    thr = round(0.9 * 256)
    rate = big.mean()
    sigma = np.sqrt(rate * (1 - rate) / big.size)
    assert abs(rate - thr / 256.0) < 5 * sigma


CASES = [
    dict(kind="dense", b=1, sq=128, sk=128, h=4, hk=4, d=64, causal=False, p=0.0),   # config 1
    dict(kind="dense", b=2, sq=37, sk=53, h=4, hk=2, d=32, causal=True, p=0.0),
    dict(kind="dense", b=1, sq=53, sk=37, h=2, hk=1, d=40, causal=True, p=0.25),      # sq>sk: empty rows
    dict(kind="dense", b=2, sq=64, sk=64, h=2, hk=2, d=128, causal=True, p=0.1),
    dict(kind="varlen", lq=[5, 64, 1, 33], lk=[7, 64, 9, 33], h=4, hk=2, d=64, causal=True, p=0.2),
    dict(kind="varlen", lq=[16, 0, 20], lk=[16, 8, 0], h=2, hk=2, d=16, causal=False, p=0.0),
]


def build(c, **kw):
    if c["kind"] == "dense":
        return cases.dense(c["b"], c["sq"], c["sk"], c["h"], c["hk"], c["d"], is_causal=c["causal"],
                           p_dropout=c["p"], seed=77, offset=3, **kw)
    return cases.varlen(c["lq"], c["lk"], c["h"], c["hk"], c["d"], is_causal=c["causal"],
                        p_dropout=c["p"], seed=77, offset=3, **kw)


def numpy_reference(pb, q, k, v, dout):
    """Loop the single-head numpy reference over (b, h) in float64."""
    qf, kf, vf = (O.from_storage(a, pb.is_bf16).astype(np.float64) for a in (q, k, v))
    dof = O.from_storage(dout, pb.is_bf16).astype(np.float64)
    out = np.zeros(pb.q_shape())
    lse = np.full((pb.batch, pb.num_heads, pb.seqlen_q), np.inf)
    dq = np.zeros(pb.q_shape()); dk = np.zeros(pb.k_shape()); dv = np.zeros(pb.k_shape())
    g = pb.num_heads // pb.num_heads_k
    for b, q0, sq, k0, sk in cases.seq_slices(pb):
        for h in range(pb.num_heads):
            hk = h // g
            args = (qf[q0:q0 + sq, h], kf[k0:k0 + sk, hk], vf[k0:k0 + sk, hk])
            kw = dict(scale=pb.softmax_scale, causal=pb.is_causal, p_dropout=pb.p_dropout, seed=pb.seed,
                      offset=pb.offset, bh=b * pb.num_heads + h)
            o, l, _, _ = R.attention_fwd(*args, **kw)
            out[q0:q0 + sq, h] = o
            lse[b, h, :sq] = l
            # the oracle's backward takes the ROUNDED forward output for D; mirror that
            a, bb, c = R.attention_bwd(*args, dof[q0:q0 + sq, h], **kw)
            dq[q0:q0 + sq, h] = a
            dk[k0:k0 + sk, hk] += bb
            dv[k0:k0 + sk, hk] += c
    return out, lse, dq, dk, dv


@pytest.mark.parametrize("c", CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items() if k != "kind"))
@pytest.mark.parametrize("is_bf16", [True, False])
def test_oracle_matches_numpy_fp64(c, is_bf16):
    pb64 = build(c, is_bf16=is_bf16, use_fp64=True)
    pb32 = build(c, is_bf16=is_bf16)
    q, k, v, dout = cases.make_inputs(pb64, seed=5)
    ro, rl, rdq, rdk, rdv = numpy_reference(pb64, q, k, v, dout)
    for pb, tol in ((pb64, 1e-9), (pb32, 2e-5)):
        f = O.forward(pb, q, k, v)
        assert cases.rel_err(f["out_f32"], ro) < max(tol, 2e-7)   # out_f32 is a float32 buffer
        fin = np.isfinite(rl)
        assert np.array_equal(np.isfinite(f["lse"]), fin)
        assert np.abs(f["lse"][fin] - rl[fin]).max() < max(tol * 10, 1e-5)
        # feed the UN-rounded output as `out` is impossible (storage only); use rounded and
        # compare with loose tolerance on D's rounding
        g = O.backward(pb, dout, q, k, v, f["out"], f["lse"])
        dtol = 2e-2 if is_bf16 else 3e-3      # D uses the storage-rounded out
        assert cases.rel_err(g["dq_f32"], rdq) < dtol
        assert cases.rel_err(g["dk_f32"], rdk) < dtol
        assert cases.rel_err(g["dv_f32"], rdv) < max(tol, 2e-6)


def test_backward_finite_differences_fp64():
    """d(sum(out*dout))/dq etc. by central differences of the numpy fp64 forward."""
    rng = np.random.default_rng(11)
    sq, sk, d = 6, 9, 8
    q, k, v, do = (rng.standard_normal(s) for s in ((sq, d), (sk, d), (sk, d), (sq, d)))
    kw = dict(scale=0.3, causal=True, p_dropout=0.3, seed=5, offset=1, bh=2)
    dq, dk, dv = R.attention_bwd(q, k, v, do, **kw)

    def loss(q_, k_, v_):
        return float((R.attention_fwd(q_, k_, v_, **kw)[0] * do).sum())

    eps = 1e-6
    for arr, grad, idx in ((q, dq, 0), (k, dk, 1), (v, dv, 2)):
        num = np.zeros_like(arr)
        for i in np.ndindex(arr.shape):
            old = arr[i]
            arr[i] = old + eps; lp = loss(q, k, v)
            arr[i] = old - eps; lm = loss(q, k, v)
            arr[i] = old
            num[i] = (lp - lm) / (2 * eps)
        assert np.abs(num - grad).max() < 1e-6, idx


def test_softmax_return_and_empty_rows():
    pb = cases.dense(1, 53, 37, 2, 1, 40, is_causal=True, p_dropout=0.25, seed=9, offset=0)
    q, k, v, _ = cases.make_inputs(pb, seed=1, with_dout=False)
    f = O.forward(pb, q, k, v, return_softmax=True)
    # rows 0..15 see no key (bottom-right alignment, sq - sk = 16): out 0, lse +inf
    assert np.all(np.isinf(f["lse"][0, :, :16])) and np.all(f["lse"][0, :, :16] > 0)
    assert np.all(O.from_storage(f["out"][:16], True) == 0)
    assert np.all(np.isfinite(f["lse"][0, :, 16:]))
    sm = O.from_storage(f["softmax"], True)[0, 0, :53, :37]
    assert np.allclose(np.abs(sm[16:]).sum(axis=1), 1.0, atol=2e-2)
    keep = O.dropout_mask(9, 0, 0.25, 0, 53, 37).astype(bool)
    nz = sm != 0
    assert np.array_equal((sm > 0)[nz], keep[nz])


def test_attn_mask_bias():
    pb = cases.dense(2, 16, 24, 2, 2, 16)
    rng = np.random.default_rng(3)
    bias = O.to_storage(rng.standard_normal((1, 2, 16, 24), dtype=np.float32), True)
    pb.attn_mask = bias
    q, k, v, dout = cases.make_inputs(pb, seed=2)
    f = O.forward(pb, q, k, v)
    qf, kf, vf = (O.from_storage(a, True).astype(np.float64) for a in (q, k, v))
    bf = O.from_storage(bias, True).astype(np.float64)
    for b in range(2):
        for h in range(2):
            o, l, _, _ = R.attention_fwd(qf[b * 16:(b + 1) * 16, h], kf[b * 24:(b + 1) * 24, h],
                                         vf[b * 24:(b + 1) * 24, h], pb.softmax_scale, bias=bf[0, h])
            assert cases.rel_err(f["out_f32"][b * 16:(b + 1) * 16, h], o) < 2e-5
            assert np.abs(f["lse"][b, h] - l).max() < 1e-4


def test_golden_vectors():
    """Committed oracle outputs (tests/golden/make_golden.py) still reproduce."""
    path = os.path.join(GOLDEN, "oracle_golden.npz")
    g = np.load(path)
    import make_golden_cases as G
    for name, pb in G.problems().items():
        q, k, v, dout = cases.make_inputs(pb, seed=G.INPUT_SEED)
        f = O.forward(pb, q, k, v)
        assert np.array_equal(f["out"], g[f"{name}.out"]), name
        assert np.allclose(f["lse"], g[f"{name}.lse"], rtol=0, atol=2e-6, equal_nan=True), name
        b = O.backward(pb, dout, q, k, v, f["out"], f["lse"])
        for t in ("dq", "dk", "dv"):
            diff = np.abs(O.from_storage(b[t], pb.is_bf16) - O.from_storage(g[f"{name}.{t}"], pb.is_bf16))
            # 1-ulp slack: libm/FMA contraction may differ between hosts
            assert cases.rel_err(O.from_storage(b[t], pb.is_bf16), O.from_storage(g[f"{name}.{t}"], pb.is_bf16)) < 1e-2, (name, t, diff.max())
