"""`return_softmax` beside the tcgen05 forward (reference flash_attn.h:15 `softmax_ptr`, :30 `return_softmax`;
flash_attn_varlen_fwd likewise): out / LSE come from the tcgen05 kernel exactly as without the flag (bit for bit) and
the S_dmask plane is written by the second pass of csrc/fa_softmax_ret.cu.  Checked per (batch, head):

  * entries the causal mask hides are exactly 0; the SIGN of every other non-zero entry is the keep bit of the written
    dropout specification (oracle.dropout_mask), bit for bit;
  * |S_dmask| equals the oracle's normalised probabilities within 1e-2 absolute (they are <= 1, 16-bit storage);
  * the plane explains the kernel's own output: (kept |S_dmask| / (1 - p)) V == out within the 2e-2 parity tolerance --
    the mask shown is the mask the shipped kernel applied.
"""
import numpy as np
import pytest

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL_REL = 2e-2


def _mask(rng, shape, neg_inf_frac=0.0):
    m = rng.standard_normal(shape, dtype=np.float32)
    if neg_inf_frac:
        m[rng.random(shape) < neg_inf_frac] = -np.inf
    return O.to_storage(m, True)


def _with_mask(pb, shape, seed=5, neg_inf_frac=0.0):
    pb.attn_mask = _mask(np.random.default_rng(seed), shape, neg_inf_frac)
    return pb


CASES = {
    "d128_causal_gqa_drop": lambda: cases.dense(2, 200, 333, 4, 2, 128, is_causal=True, p_dropout=0.3,
                                                seed=2**40 + 3, offset=2**33 + 9),
    "d64_fp16_full": lambda: cases.dense(1, 130, 70, 2, 2, 64, is_bf16=False),
    "d72_drop": lambda: cases.dense(1, 65, 129, 2, 1, 72, p_dropout=0.5, seed=7, offset=1),
    "d256_causal": lambda: cases.dense(1, 96, 160, 2, 2, 256, is_causal=True, p_dropout=0.1, seed=3),
    "rows_without_keys": lambda: cases.dense(1, 150, 60, 2, 1, 64, is_causal=True, p_dropout=0.25, seed=9),
    "packed_causal_drop": lambda: cases.varlen([5, 200, 1, 129], [7, 200, 64, 300], 4, 2, 128, is_causal=True,
                                               p_dropout=0.2, seed=0xDEADBEEFCAFEF00D, offset=0xFFFFFFFF00000001),
    "packed_empty_seq": lambda: cases.varlen([16, 0, 70], [16, 8, 0], 2, 2, 64),
    "bias_drop": lambda: _with_mask(cases.dense(2, 100, 140, 2, 2, 128, p_dropout=0.1, seed=11), (1, 2, 100, 140)),
    "bias_minus_inf_causal": lambda: _with_mask(cases.dense(1, 128, 128, 2, 1, 64, is_causal=True), (1, 1, 128, 128),
                                                neg_inf_frac=0.3),
    "batch_base": lambda: cases.dense(2, 64, 64, 2, 2, 64, p_dropout=0.4, seed=5, offset=2**32, batch_base=7),
}


def check_plane(name, pb, q, k, v, recon=True):
    """Runs the forward with and without `return_softmax` and checks the plane (module docstring)."""
    import gpu_util as G
    from flash_attention_hip_b200 import api
    plain = G.run_fwd(pb, q, k, v)
    got = G.run_fwd(pb, q, k, v, return_softmax=True)
    if plain["path"] != api.PATH_TCGEN05:        # (softmax_scale <= 0 and forced paths: the all-shapes kernel writes it itself)
        assert got["path"] == plain["path"]
    else:
        assert got["path"] == api.PATH_TCGEN05
        assert api.last_launch_count() == 3      # rng_state, tc_fwd_kernel, tc_softmax_ret_kernel
    assert np.array_equal(got["out"], plain["out"])
    assert np.array_equal(got["lse"].view(np.uint32), plain["lse"].view(np.uint32))

    ref = O.forward(pb, q, k, v, return_softmax=True)
    sm = O.from_storage(got["softmax"], pb.is_bf16)
    rs = O.from_storage(ref["softmax"], pb.is_bf16)
    assert sm.shape == rs.shape
    out = G.f32(got["out"], pb)
    vf = O.from_storage(v, pb.is_bf16)
    group = pb.num_heads // pb.num_heads_k
    scale_keep = 0.0 if pb.p_dropout >= 1 else 1.0 / (1.0 - pb.p_dropout)
    for b, q0, sq, k0, sk in cases.seq_slices(pb):
        # nothing outside [0, sq) x [0, sk) is touched (the wrapper hands over a zeroed plane)
        assert not sm[b, :, sq:, :].any() and not sm[b, :, :, sk:].any()
        if sq == 0 or sk == 0:
            continue
        i = np.arange(sq)[:, None]
        j = np.arange(sk)[None, :]
        vis = (j <= i + (sk - sq)) if pb.is_causal else np.ones((sq, sk), bool)
        for h in range(pb.num_heads):
            s = sm[b, h, :sq, :sk]
            assert not s[~vis].any(), f"{name}: hidden entries of ({b}, {h}) are not zero"
            nz = s != 0
            if pb.p_dropout > 0:
                keep = O.dropout_mask(pb.seed, pb.offset, pb.p_dropout, (pb.batch_base + b) * pb.num_heads + h,
                                      sq, sk).astype(bool)
                assert np.array_equal((s > 0)[nz], keep[nz]), f"{name}: keep bits of ({b}, {h})"
            else:
                assert (s >= 0).all()
            assert np.abs(np.abs(s) - np.abs(rs[b, h, :sq, :sk])).max() < 1e-2, f"{name}: values of ({b}, {h})"
            if not recon:
                continue
            kept = np.where(s > 0, s, 0.0).astype(np.float64) * scale_keep
            recon_o = kept @ vf[k0:k0 + sk, h // group, :].astype(np.float64)
            mine = out[q0:q0 + sq, h, :]
            if np.abs(recon_o).max() == 0:
                assert not mine.any()
            else:
                assert cases.rel_err(mine, recon_o) < TOL_REL, f"{name}: out of ({b}, {h})"


@pytest.mark.parametrize("name", list(CASES))
def test_return_softmax_on_the_tcgen05_forward(name):
    pb = CASES[name]()
    q, k, v, _ = cases.make_inputs(pb, seed=21, with_dout=False)
    check_plane(name, pb, q, k, v)


@pytest.mark.parametrize("i", range(0, 160, 4))
def test_return_softmax_random_problems(i):
    """Every fourth problem of the random sweep (tests/test_gpu_random_sweep.py: head sizes 8..256, ragged lengths, GQA,
    dtype, causal, dropout, attn_mask with -inf entries, dense or packed) with `return_softmax`."""
    from test_gpu_random_sweep import _draw
    pb, _ = _draw(i)
    q, k, v, _ = cases.make_inputs(pb, seed=50 + i, with_dout=False)
    check_plane(f"sweep/{i}", pb, q, k, v, recon=False)


def test_return_softmax_needs_the_plane():
    import torch
    from flash_attention_hip_b200 import api
    lib = api.lib()
    q = torch.zeros((1, 128, 2, 64), dtype=torch.bfloat16, device="cuda")
    out = torch.empty_like(q)
    lse = torch.empty((1, 2, 128), dtype=torch.float32, device="cuda")
    ok = lib.flash_attn_fwd(q.data_ptr(), q.data_ptr(), q.data_ptr(), None, out.data_ptr(), None, lse.data_ptr(),
                            1, 128, 128, 128, 128, 2, 2, 64, 64, 0.0, 0.125, 8.0, False, True, True, None, 0, 0,
                            None, None)
    assert not ok and b"softmax_ptr is NULL" in lib.flash_attn_error()
