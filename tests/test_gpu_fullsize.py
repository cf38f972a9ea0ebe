"""Parity at BASELINE.json's FULL sizes (configs 2, 3 and 4), where the CPU oracle would take minutes:
  * sampled (batch, head) units of the full-size call against fp32 math on the same device, with the
    dropout mask taken from the oracle's Philox (oracle.dropout_mask) -- north_star tolerances;
  * size-independent properties: causality (later keys cannot change earlier rows, bit for bit),
    softmax normalisation (V = 1 gives out = 1), dropout keep rate, run-to-run equality of the forward.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

TOL_REL, TOL_LSE = 2e-2, 1e-3


def _rand(shape, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(shape, generator=g, device="cuda", dtype=torch.float32).bfloat16()


def _rel(a, ref):
    return float((a.float() - ref).abs().max() / ref.abs().max())


def _unit_reference(q, k, v, do, scale, causal, keep, rp, out_given=None):
    """fp32 attention forward + backward of ONE (batch, head) unit; keep: bool [sq, sk] or None.
    out_given: the `out` tensor the backward entry point was handed (reference flash_attn.h:72): softmax_d =
    rowsum(dO o out) is defined on THAT tensor (16-bit), not on an unrounded recomputation."""
    q, k, v, do = (t.float() for t in (q, k, v, do))
    s = (q @ k.t()) * scale
    if causal:
        sq, sk = s.shape
        s = s.masked_fill(torch.ones(sq, sk, dtype=torch.bool, device=s.device).triu(1 + sk - sq), float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.exp(s - lse[:, None])
    pd = p if keep is None else p * keep * rp
    out = pd @ v
    dv = pd.t() @ do
    dp = do @ v.t()
    if keep is not None:
        dp = dp * keep * rp
    d = (do * (out if out_given is None else out_given.float())).sum(-1, keepdim=True)
    ds = p * (dp - d)
    return out, lse, scale * (ds @ k), scale * (ds.t() @ q), dv


@pytest.mark.parametrize("cfg", ["c3_fwd_bwd_dropout", "c2_fwd"])
def test_full_size_sampled_units(cfg):
    import parity
    from flash_attention_hip_b200 import api
    if cfg == "c3_fwd_bwd_dropout":      # BASELINE.json configs[2]
        b, h, s, d, p_drop, seed, offset, units = 16, 32, 4096, 128, 0.1, 1234, 0, [(0, 0), (7, 13), (15, 31)]
    else:                                # configs[1]
        b, h, s, d, p_drop, seed, offset, units = 8, 16, 2048, 128, 0.0, 0, 0, [(0, 0), (3, 9), (7, 15)]
    q, k, v, do = (_rand((b, s, h, d), 100 + i) for i in range(4))
    scale = d ** -0.5
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, p_dropout=p_drop, is_causal=True, seed=seed, offset=offset)
    assert api.last_path() == api.PATH_TCGEN05
    grads = None
    if cfg.startswith("c3"):
        grads = api.flash_attn_bwd(do, q, k, v, out, lse, p_dropout=p_drop, is_causal=True, seed=seed, offset=offset)
        assert api.last_path() == api.PATH_TCGEN05
    for bi, hi in units:
        keep, rp = None, 1.0
        if p_drop > 0:
            keep = torch.from_numpy(O.dropout_mask(seed, offset, p_drop, bi * h + hi, s, s).astype(np.bool_)).cuda()
            rp = 1.0 / (1.0 - p_drop)
        args = (q[bi, :, hi], k[bi, :, hi], v[bi, :, hi], do[bi, :, hi], scale, True, keep, rp, out[bi, :, hi])
        r_out, r_lse, r_dq, r_dk, r_dv = _unit_reference(*args)
        l_out, l_dq, l_dk, l_dv = _lowprec_unit(*args)
        tag = f"fullsize/{cfg}/b{bi}h{hi}"
        n = lambda t: t.float().cpu().numpy()
        parity.check(tag, "out", n(out[bi, :, hi]), n(r_out), n(l_out))
        parity.check_lse(tag, n(lse[bi, hi]), n(r_lse))
        if grads is not None:
            parity.check(tag, "dq", n(grads[0][bi, :, hi]), n(r_dq), n(l_dq))
            parity.check(tag, "dk", n(grads[1][bi, :, hi]), n(r_dk), n(l_dk))
            parity.check(tag, "dv", n(grads[2][bi, :, hi]), n(r_dv), n(l_dv))


def test_full_size_properties():
    from flash_attention_hip_b200 import api
    b, h, s, d = 4, 32, 4096, 128
    q, k, v = (_rand((b, s, h, d), 200 + i) for i in range(3))
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=True)
    # run-to-run equality
    out2, lse2, _, _ = api.flash_attn_fwd(q, k, v, is_causal=True)
    assert torch.equal(out, out2) and torch.equal(lse, lse2)
    # causality: keys/values at positions >= t cannot change rows < t, bit for bit
    t = 2560                                  # tile-aligned: rows below t never touch a tile that changed
    k2, v2 = k.clone(), v.clone()
    k2[:, t:] = _rand((b, s - t, h, d), 301)
    v2[:, t:] = _rand((b, s - t, h, d), 302)
    out3, lse3, _, _ = api.flash_attn_fwd(q, k2, v2, is_causal=True)
    assert torch.equal(out[:, :t], out3[:, :t]) and torch.equal(lse[:, :, :t], lse3[:, :, :t])
    assert not torch.equal(out[:, t:], out3[:, t:])
    # normalisation: V = 1 -> every output is 1 (P rows sum to 1 after bf16 rounding of P)
    ones = torch.ones_like(v)
    o1, _, _, _ = api.flash_attn_fwd(q, k, ones, is_causal=True)
    assert float((o1.float() - 1).abs().max()) < 1e-2
    # dropout keep rate with V = 1: E[out] = 1, and the mask depends on (seed, offset) only
    od, _, _, _ = api.flash_attn_fwd(q, k, ones, is_causal=True, p_dropout=0.3, seed=5, offset=7)
    od2, _, _, _ = api.flash_attn_fwd(q, k, ones, is_causal=True, p_dropout=0.3, seed=5, offset=7)
    oe, _, _, _ = api.flash_attn_fwd(q, k, ones, is_causal=True, p_dropout=0.3, seed=5, offset=8)
    assert torch.equal(od, od2) and not torch.equal(od, oe)
    assert abs(float(od[:, 2048:].float().mean()) - 1.0) < 5e-3


# ---- BASELINE.json configs[3] exactly as SURVEY.md section 8(d) words it -------------------------------------------
C4_LENS = [128, 8192, 512, 4096, 1000, 2048, 256, 6000]     # 22232 tokens, includes non-multiples of 128
C4_H, C4_HK = 32, 8


def _lowprec_unit(q, k, v, do, scale, causal, keep, rp, out_given=None):
    """The same unit in a 16-bit pipeline (scores / P / dS / outputs rounded to bf16 as they are produced):
    its error against the fp32 unit calibrates the second parity criterion (tests/parity.py)."""
    bf = torch.bfloat16
    s = ((q @ k.t()).float() * scale)
    if causal:
        sq, sk = s.shape
        s = s.masked_fill(torch.ones(sq, sk, dtype=torch.bool, device=s.device).triu(1 + sk - sq), float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.exp(s - lse[:, None])
    pd = (p if keep is None else p * keep * rp).to(bf)
    out = (pd @ v).to(bf)
    dv = (pd.t() @ do)
    dp = (do @ v.t()).float()
    if keep is not None:
        dp = dp * keep * rp
    d = (do.float() * (out if out_given is None else out_given).float()).sum(-1, keepdim=True)
    ds = (p * (dp - d)).to(bf)
    return out.float(), (scale * (ds @ k).float()), (scale * (ds.t() @ q).float()), dv.float()


@pytest.mark.parametrize("d,causal,p_drop", [(128, True, 0.0), (128, False, 0.0), (64, True, 0.0), (64, False, 0.0),
                                             (128, True, 0.1), (64, True, 0.1)])
def test_config4_full_mix_sampled_units(d, causal, p_drop):
    """All eight lengths, H = 32 / 8 (GQA), D = 64 and 128, causal and not, forward + backward, through the
    packed entry points.  One kv head per sequence is checked in full: its four query heads against fp32 math
    on the device (dropout mask from the oracle's Philox), dK / dV against the sum over the group."""
    import itertools
    import parity
    from flash_attention_hip_b200 import api
    lens = C4_LENS
    cu = torch.tensor([0] + list(itertools.accumulate(lens)), dtype=torch.int32, device="cuda")
    total, mx = sum(lens), max(lens)
    q, do = _rand((total, C4_H, d), 400), _rand((total, C4_H, d), 401)
    k, v = _rand((total, C4_HK, d), 402), _rand((total, C4_HK, d), 403)
    seed, offset = 3, 2**33 + 5
    kw = dict(p_dropout=p_drop, is_causal=causal, seed=seed, offset=offset)
    out, lse, _, _ = api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, **kw)
    assert api.last_path() == api.PATH_TCGEN05
    dq, dk, dv = api.flash_attn_varlen_bwd(do, q, k, v, out, lse, cu, cu, mx, mx, **kw)
    assert api.last_path() == api.PATH_TCGEN05
    scale, group, rp = d ** -0.5, C4_H // C4_HK, 1.0 / (1.0 - p_drop)
    label = f"c4_full/d{d}_{'causal' if causal else 'full'}_p{p_drop}"
    for b, n in enumerate(lens):
        if p_drop > 0 and n > 2048:
            continue                       # the oracle's mask of an 8192 x 8192 unit is minutes of CPU; five sequences carry it
        t0 = int(cu[b])
        hk = (3 * b + 1) % C4_HK           # a different kv head per sequence
        sl = slice(t0, t0 + n)
        rdk = torch.zeros((n, d), device="cuda"); rdv = torch.zeros((n, d), device="cuda")
        ldk = torch.zeros((n, d), device="cuda"); ldv = torch.zeros((n, d), device="cuda")
        for h in range(hk * group, (hk + 1) * group):
            keep = None
            if p_drop > 0:
                keep = torch.from_numpy(O.dropout_mask(seed, offset, p_drop, b * C4_H + h, n, n).astype(np.bool_)).cuda()
            args = (q[sl, h], k[sl, hk], v[sl, hk], do[sl, h], scale, causal, keep, rp, out[sl, h])
            r_out, r_lse, r_dq, r_dk, r_dv = _unit_reference(*args)
            l_out, l_dq, l_dk, l_dv = _lowprec_unit(*args)
            tag = f"{label}/seq{b}(n={n})/h{h}"
            parity.check(tag, "out", out[sl, h].float().cpu().numpy(), r_out.cpu().numpy(), l_out.cpu().numpy())
            parity.check_lse(tag, lse[b, h, :n].cpu().numpy(), r_lse.cpu().numpy())
            parity.check(tag, "dq", dq[sl, h].float().cpu().numpy(), r_dq.cpu().numpy(), l_dq.cpu().numpy())
            rdk += r_dk; rdv += r_dv; ldk += l_dk; ldv += l_dv
        tag = f"{label}/seq{b}(n={n})/hk{hk}"
        parity.check(tag, "dk", dk[sl, hk].float().cpu().numpy(), rdk.cpu().numpy(), ldk.bfloat16().float().cpu().numpy())
        parity.check(tag, "dv", dv[sl, hk].float().cpu().numpy(), rdv.cpu().numpy(), ldv.bfloat16().float().cpu().numpy())


@pytest.mark.parametrize("d,hk", [(128, 1), (64, 2), (256, 2)])
def test_return_softmax_long_sequence(d, hk):
    """`return_softmax` at S = 4096 (32 key tiles: the K ring and the two S accumulators of csrc/fa_softmax_ret.cu wrap many
    times): every head's plane against fp32 math on the device (1e-2 absolute), hidden entries exactly zero, signs against
    the oracle's Philox for one head, and out / LSE bit-equal to the call without the flag."""
    from flash_attention_hip_b200 import api
    b, h, s = 1, 2, 4096
    q = _rand((b, s, h, d), 31)
    k, v = _rand((b, s, hk, d), 32), _rand((b, s, hk, d), 33)
    kw = dict(p_dropout=0.2, is_causal=True, seed=2**40 + 5, offset=2**33 + 1)
    out0, lse0, _, _ = api.flash_attn_fwd(q, k, v, **kw)
    out, lse, sm, _ = api.flash_attn_fwd(q, k, v, return_softmax=True, **kw)
    assert api.last_path() == api.PATH_TCGEN05
    assert torch.equal(out, out0) and torch.equal(lse.view(torch.int32), lse0.view(torch.int32))
    hidden = torch.ones(s, s, dtype=torch.bool, device="cuda").triu(1)
    for hh in range(h):
        sc = (q[0, :, hh].float() @ k[0, :, hh // (h // hk)].float().t()) * d ** -0.5
        p_ref = torch.exp(sc.masked_fill(hidden, float("-inf")) - lse[0, hh][:, None])
        plane = sm[0, hh, :s, :s].float()
        assert not plane[hidden].any()
        assert float((plane.abs() - p_ref).abs().max()) < 1e-2
        # kept |P| / (1 - p) times V reproduces the kernel's own output
        recon = (plane.clamp_min(0) / 0.8) @ v[0, :, hh // (h // hk)].float()
        assert _rel(out[0, :, hh], recon) < TOL_REL
    keep = torch.from_numpy(O.dropout_mask(kw["seed"], kw["offset"], 0.2, 1, s, s).astype(bool)).cuda()   # head 1 of batch 0
    plane = sm[0, 1, :s, :s].float()
    nz = plane != 0
    assert torch.equal((plane > 0)[nz], keep[nz])
    frac = float((plane[~hidden] > 0).float().mean())
    assert abs(frac - 205 / 256) < 5e-3            # threshold round(0.8 * 256) = 205
