"""GPU tests of the multi-GPU pieces that can run on ONE device (the N>1 schedules are covered by
tests/test_multi_rank_cpu.py under gloo and by tests/cuda/ring_check.py under torchrun):
  * batch_base: two shard calls reproduce the unsharded call bit-exactly, dropout included,
    and agree with the oracle;
  * cp_merge / cp_accumulate / cp_store (csrc/fa_cp.cu) against torch fp32;
  * the ring driver at world_size 1 (gloo group of one rank) equals the plain entry points.
"""
import os

import numpy as np
import pytest
import torch

import cases
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _rand(shape, seed, dtype=torch.bfloat16):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return torch.randn(shape, generator=g, device="cuda", dtype=torch.float32).to(dtype)


@pytest.mark.parametrize("d,s", [(128, 384), (64, 200), (96, 130)])   # all three ride the tcgen05 kernels (96 padded to 128)
def test_batch_shards_equal_unsharded_call(d, s):
    from flash_attention_hip_b200 import api
    from flash_attention_hip_b200.shard import BatchShardedAttention
    gb, h, hk = 3, 4, 2
    q, k, v, do = _rand((gb, s, h, d), 1), _rand((gb, s, hk, d), 2), _rand((gb, s, hk, d), 3), _rand((gb, s, h, d), 4)
    kw = dict(p_dropout=0.2, is_causal=True, seed=99, offset=3)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, **kw)
    dq, dk, dv = api.flash_attn_bwd(do, q, k, v, out, lse, **kw)
    for rank in range(2):
        sh = BatchShardedAttention(gb, rank, 2)
        ql, kl, vl, dol = (sh.local_slice(x).contiguous() for x in (q, k, v, do))
        o2, l2 = sh.fwd(ql, kl, vl, causal=True, p_dropout=0.2, seed=99, offset=3)
        assert torch.equal(o2, sh.local_slice(out)) and torch.equal(l2, sh.local_slice(lse))
        # the tcgen05 backward adds dQ tiles in arrival order: compare within a tight tolerance
        g2 = sh.bwd(dol, ql, kl, vl, o2, l2, causal=True, p_dropout=0.2, seed=99, offset=3)
        for got, ref in zip(g2, (dq, dk, dv)):
            ref = sh.local_slice(ref).float()
            assert float((got.float() - ref).abs().max() / ref.abs().max()) < 2e-3
    # and against the oracle, which holds the same batch_base rule (rank 1 = global batch 2)
    import gpu_util as G
    pb = cases.dense(1, s, s, h, hk, d, is_causal=True, p_dropout=0.2, seed=99, offset=3)
    pb.batch_base = 2
    ref = O.forward(pb, G.to_bits(q[2:]).reshape(pb.q_shape()), G.to_bits(k[2:]).reshape(pb.k_shape()),
                    G.to_bits(v[2:]).reshape(pb.k_shape()))
    err = np.abs(G.f32(G.to_bits(out[2:]).reshape(pb.q_shape()), pb) - ref["out_f32"]).max() / np.abs(ref["out_f32"]).max()
    assert err < 2e-2


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_cp_kernels_match_torch(dtype):
    from flash_attention_hip_b200 import api
    rows, total, h, d = 333, 500, 3, 128
    o_acc = torch.randn(total, h, d, device="cuda")
    lse_acc = torch.randn(h, total, device="cuda") * 3
    lse_acc[:, 100:150] = float("inf")                         # rows that have seen nothing yet
    o_acc[100:150] = 0
    o_blk = torch.randn(rows, h, d, device="cuda").to(dtype)
    lse_blk = torch.randn(h, rows, device="cuda") * 3
    lse_blk[:, 40:60] = float("inf")                           # this step saw no key for them
    r0 = 90
    ref_o, ref_l = o_acc.clone(), lse_acc.clone()
    la, lb = ref_l[:, r0:r0 + rows], lse_blk
    la_f = torch.where(torch.isinf(la), torch.full_like(la, -float("inf")), la)
    lb_f = torch.where(torch.isinf(lb), torch.full_like(lb, -float("inf")), lb)
    ln = torch.logaddexp(la_f, lb_f)
    wa = torch.nan_to_num(torch.exp(la_f - ln), nan=1.0)
    wb = torch.nan_to_num(torch.exp(lb_f - ln), nan=0.0)
    ref_o[r0:r0 + rows] = wa.t().unsqueeze(-1) * ref_o[r0:r0 + rows] + wb.t().unsqueeze(-1) * o_blk.float()
    ref_l[:, r0:r0 + rows] = torch.where(torch.isinf(ln), torch.full_like(ln, float("inf")), ln)
    api.cp_merge(o_acc[r0:r0 + rows], lse_acc[:, r0:r0 + rows], o_blk, lse_blk)
    torch.cuda.synchronize()
    assert torch.allclose(o_acc, ref_o, rtol=1e-5, atol=1e-5)
    assert torch.allclose(lse_acc, ref_l, rtol=1e-5, atol=1e-5)
    # both empty stays empty
    assert bool(torch.isinf(lse_acc[:, 130:150]).all()) and float(o_acc[130:150].abs().max()) == 0.0

    acc = torch.randn(1000, 7, device="cuda")
    blk = torch.randn(1000, 7, device="cuda").to(dtype)
    ref = acc + blk.float()
    api.cp_accumulate(acc, blk)
    assert torch.equal(acc, ref)
    dst = torch.empty(1000, 7, device="cuda", dtype=dtype)
    api.cp_store(dst, acc)
    assert torch.equal(dst, acc.to(dtype))
    # a length that is not a multiple of 8 (vector body + scalar tail), and buffers that are not 16-byte aligned (scalar kernels)
    for n, off in ((3003, 0), (3003, 1), (5, 0)):
        acc = torch.randn(n + off, device="cuda")[off:]
        blk = torch.randn(n + off, device="cuda").to(dtype)[off:]
        ref = acc + blk.float()
        api.cp_accumulate(acc, blk)
        assert torch.equal(acc, ref), (n, off)
        dst = torch.empty(n + off, device="cuda", dtype=dtype)[off:]
        api.cp_store(dst, acc)
        assert torch.equal(dst, acc.to(dtype)), (n, off)


def test_ring_world1_equals_plain_calls():
    import torch.distributed as dist
    from flash_attention_hip_b200 import api, ring
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29531")
    dist.init_process_group("gloo", rank=0, world_size=1)
    try:
        b, s, h, hk, d = 2, 512, 4, 2, 128
        q, k, v, do = _rand((b, s, h, d), 5), _rand((b, s, hk, d), 6), _rand((b, s, hk, d), 7), _rand((b, s, h, d), 8)
        for causal in (True, False):
            out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=causal)
            dq, dk, dv = api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=causal)
            o2, l2 = ring.ring_attn_fwd(q, k, v, causal=causal)
            assert torch.equal(o2, out) and torch.allclose(l2, lse, atol=1e-6)
            g2 = ring.ring_attn_bwd(do, q, k, v, o2, l2, causal=causal)
            for got, ref in zip(g2, (dq, dk, dv)):
                assert float((got.float() - ref.float()).abs().max() / ref.float().abs().max()) < 2e-3
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_ring_c_entry_points_multi_gpu(world):
    """flash_attn_b200_ring_fwd / _bwd (csrc/fa_ring.cu) over `world` GPUs of this box against a single-GPU
    full-sequence run (tests/cuda/ring_check.py: 2e-2 relative on out / dq / dk / dv, 1e-3 on LSE, every rank).
    Skipped where fewer GPUs are visible; bench.py's `ring_c5` key repeats the check inside the scaling run."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29540 + world), os.path.join(root, "tests", "cuda", "ring_check.py"), "8192", "1", "notime"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("parity OK") == world, r.stdout[-3000:]


def test_ring_dropout_multi_gpu():
    """Dropout in ring mode (flash_attn_b200_ring_fwd_dropout / _bwd_dropout) over 2 GPUs: every rank's rows against the
    host-side restatement of the written per-step masks (tests/ring_ref.py) and against the Python schedule
    (tests/cuda/ring_dropout_check.py).  Skipped where fewer GPUs are visible; the gloo test
    tests/test_multi_rank_cpu.py checks the same specification on CPU."""
    import subprocess
    import sys
    import torch
    world = 2
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29547", os.path.join(root, "tests", "cuda", "ring_dropout_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("dropout parity OK") == 2 * world, r.stdout[-3000:]


def test_ring_dropout_world1_equals_plain_call_with_the_step_offset():
    """World size 1: the one step of the ring IS the plain call with step_offset(offset, 0, 0, 1) = offset and
    batch_base = the batch entry -- the dropout entry points against the ordinary ones, bit for bit on the forward."""
    import torch.distributed as dist
    from flash_attention_hip_b200 import api, ring
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("gloo", rank=0, world_size=1)
    try:
        b, s, h, hk, d = 3, 384, 4, 2, 128
        q, k, v, do = _rand((b, s, h, d), 15), _rand((b, s, hk, d), 16), _rand((b, s, hk, d), 17), _rand((b, s, h, d), 18)
        kw = dict(p_dropout=0.25, seed=2**40 + 1, offset=2**33 + 2)
        for causal in (True, False):
            out, lse, _, _ = api.flash_attn_fwd(q, k, v, is_causal=causal, **kw)
            dq, dk, dv = api.flash_attn_bwd(do, q, k, v, out, lse, is_causal=causal, **kw)
            o2, l2 = ring.ring_attn_fwd(q, k, v, causal=causal, **kw)
            assert torch.equal(o2, out) and torch.allclose(l2, lse, atol=1e-6)
            g2 = ring.ring_attn_bwd(do, q, k, v, o2, l2, causal=causal, **kw)
            for got, ref in zip(g2, (dq, dk, dv)):
                assert float((got.float() - ref.float()).abs().max() / ref.float().abs().max()) < 2e-3
            o0, _ = ring.ring_attn_fwd(q, k, v, causal=causal)
            assert not torch.equal(o0, out)          # and it is a different result from the call without dropout
    finally:
        ring.close_comms()
        dist.destroy_process_group()
