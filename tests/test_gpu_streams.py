"""Two streams of one device driven at the same time (a framework with a side stream; reference flash_attn.h:32 `stream`):
every entry point enqueues on the stream it is given and shares nothing with a concurrent call but the library's own
device-side pools (item counters of the persistent kernels, per-SM tensor-map slots of the packed layout), which are built
for exactly this (csrc/fa_common.cuh `SchedSlot`, csrc/ptx_sm100.cuh `TensorMapSlot`).  Calls of two different problems
are interleaved on two streams without any synchronisation between them; every result must equal the one the same
call gave alone -- bit for bit (forward; backward with num_splits = 1), within 1e-2 where dQ adds arrive in any order."""
import pytest

pytestmark = pytest.mark.gpu


def _dense_problem(torch, api, dev_gen, b, s, h, hk, d, deterministic, **kw):
    q = torch.randn(b, s, h, d, generator=dev_gen, device="cuda").bfloat16()
    k = torch.randn(b, s, hk, d, generator=dev_gen, device="cuda").bfloat16()
    v = torch.randn(b, s, hk, d, generator=dev_gen, device="cuda").bfloat16()
    do = torch.randn(b, s, h, d, generator=dev_gen, device="cuda").bfloat16()
    sr, dr = api.round_up(s, 128), api.round_up(d, 32)
    st = dict(out=torch.empty_like(q), lse=torch.empty((b, h, s), dtype=torch.float32, device="cuda"),
              grads=(torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)),
              scratch=(torch.empty((b, h, sr), dtype=torch.float32, device="cuda"),
                       torch.empty((b, sr, h, dr), dtype=torch.float32, device="cuda")))

    def run(stream):
        api.flash_attn_fwd(q, k, v, out=st["out"], lse=st["lse"], stream=stream, **kw)
        api.flash_attn_bwd(do, q, k, v, st["out"], st["lse"], scratch=st["scratch"], grads=st["grads"],
                           deterministic=deterministic, stream=stream, **kw)
    return st, run


def _packed_problem(torch, api, dev_gen, lens, h, hk, d, deterministic, **kw):
    total, mx, nb = sum(lens), max(lens), len(lens)
    cu = torch.tensor([0] + list(__import__("itertools").accumulate(lens)), dtype=torch.int32, device="cuda")
    q = torch.randn(total, h, d, generator=dev_gen, device="cuda").bfloat16()
    k = torch.randn(total, hk, d, generator=dev_gen, device="cuda").bfloat16()
    v = torch.randn(total, hk, d, generator=dev_gen, device="cuda").bfloat16()
    do = torch.randn(total, h, d, generator=dev_gen, device="cuda").bfloat16()
    sr, dr = api.round_up(mx, 128), api.round_up(d, 32)
    st = dict(out=torch.empty_like(q), lse=torch.full((nb, h, mx), float("inf"), dtype=torch.float32, device="cuda"),
              grads=(torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)),
              scratch=(torch.empty((nb, h, sr), dtype=torch.float32, device="cuda"),
                       torch.empty((nb, sr, h, dr), dtype=torch.float32, device="cuda")))

    def run(stream):
        api.flash_attn_varlen_fwd(q, k, v, cu, cu, mx, mx, out=st["out"], lse=st["lse"], stream=stream, **kw)
        api.flash_attn_varlen_bwd(do, q, k, v, st["out"], st["lse"], cu, cu, mx, mx, scratch=st["scratch"],
                                  grads=st["grads"], deterministic=deterministic, stream=stream, **kw)
    return st, run


def _snapshot(st):
    return [st["out"].clone(), st["lse"].clone()] + [g.clone() for g in st["grads"]]


def _wipe(st):
    st["out"].zero_()
    for g in st["grads"]:
        g.zero_()


@pytest.mark.parametrize("deterministic", [True, False])
def test_interleaved_calls_on_two_streams(deterministic):
    import torch
    from flash_attention_hip_b200 import api
    gen = torch.Generator(device="cuda").manual_seed(17)
    # both problems have more work items than SMs (the persistent path) and differ in kernel build (head size, layout)
    a_st, a_run = _dense_problem(torch, api, gen, 8, 1024, 16, 4, 128, deterministic,
                                 p_dropout=0.1, is_causal=True, seed=3, offset=2**33 + 1)
    b_st, b_run = _packed_problem(torch, api, gen, [700, 128, 1500, 33, 1024, 257, 900, 2048], 8, 8, 64, deterministic,
                                  p_dropout=0.2, is_causal=True, seed=2**40 + 9, offset=5)
    a_run(None); b_run(None)
    torch.cuda.synchronize()
    a_ref, b_ref = _snapshot(a_st), _snapshot(b_st)

    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    for rounds in range(2):
        _wipe(a_st); _wipe(b_st)
        torch.cuda.synchronize()
        for _ in range(6):                      # no synchronisation between the two streams: their kernels overlap
            a_run(sa)
            b_run(sb)
        torch.cuda.synchronize()
        for name, st, ref in (("dense", a_st, a_ref), ("packed", b_st, b_ref)):
            got = _snapshot(st)
            assert torch.equal(got[0], ref[0]), f"{name}: out"
            assert torch.equal(got[1].view(torch.int32), ref[1].view(torch.int32)), f"{name}: lse"
            for t, (x, r) in zip(("dq", "dk", "dv"), zip(got[2:], ref[2:])):
                if deterministic:
                    assert torch.equal(x, r), f"{name}: {t}"
                else:
                    err = (x.float() - r.float()).abs().max().item() / r.float().abs().max().item()
                    assert err < 1e-2, (name, t, err)
