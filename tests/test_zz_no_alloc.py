"""flash_attn_b200_bwd_static_scratch (include/flash_attn_b200_ext.h): the dense backward whose scratch is the
library's static device array, for callers that must not allocate.  Checked through the C ABI against the ordinary
flash_attn_bwd on two calls in a row (the second finds the first call's leftovers in the shared array), and against
its stated size limit.  Collected last on purpose (file name)."""
import numpy as np
import pytest
import torch

import cases
import gpu_util as G
from flash_attention_hip_b200 import api
from oracle import oracle as O

pytestmark = pytest.mark.gpu

# same kernels, same inputs: only the arrival order of the fp32 dQ reduce-adds differs between two launches
TOL = 1e-2


def _one_call(pb, seed):
    q, k, v, dout = cases.make_inputs(pb, seed=seed)
    ref = O.forward(pb, q, k, v)
    shp_q = (pb.batch, pb.seqlen_q, pb.num_heads, pb.head_size)
    shp_k = (pb.batch, pb.seqlen_k, pb.num_heads_k, pb.head_size)
    t = [G.to_dev(a, pb.is_bf16) for a in (dout, q, k, v, ref["out"])]
    t = [t[0].view(shp_q), t[1].view(shp_q), t[2].view(shp_k), t[3].view(shp_k), t[4].view(shp_q)]
    lse = torch.from_numpy(np.ascontiguousarray(ref["lse"], dtype=np.float32)).cuda()
    kw = dict(p_dropout=pb.p_dropout, softmax_scale=pb.softmax_scale, is_causal=pb.is_causal, seed=pb.seed,
              offset=pb.offset)
    want = api.flash_attn_bwd(*t, lse, **kw)
    want_path = api.last_path()
    torch.cuda.synchronize()
    g = G.Guarded()
    grads = (g.make(shp_q, t[1].dtype), g.make(shp_k, t[2].dtype), g.make(shp_k, t[3].dtype))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    got = api.flash_attn_bwd_static_scratch(*t, lse, grads=grads, **kw)
    assert torch.cuda.memory_allocated() == before
    assert api.last_path() == want_path == api.PATH_TCGEN05
    torch.cuda.synchronize()
    g.check()
    # and both against the oracle, so that agreeing on garbage cannot pass
    rb = O.backward(pb, dout, q, k, v, ref["out"], ref["lse"])
    for name, a, b in zip(("dq", "dk", "dv"), got, want):
        fa, fb = a.float().cpu().numpy(), b.float().cpu().numpy()
        assert np.isfinite(fa).all(), name
        assert cases.rel_err(fa, fb) < TOL, (name, cases.rel_err(fa, fb))
        assert cases.rel_err(fa.reshape(rb[name + "_f32"].shape), rb[name + "_f32"]) < 2e-2, name


def test_static_scratch_backward_matches_flash_attn_bwd_twice_in_a_row():
    # first call: 9.1 MiB of the array, dropout + causal, GQA
    _one_call(cases.dense(2, 1024, 1024, 8, 2, 128, is_causal=True, p_dropout=0.1, seed=77, offset=5), seed=11)
    # second call: another shape and other inputs over the first call's leftovers (ragged tiles, D=64, fp16)
    _one_call(cases.dense(3, 333, 450, 4, 4, 64, is_causal=False, p_dropout=0.0, is_bf16=False), seed=12)
    # third: the first problem again, other inputs
    _one_call(cases.dense(2, 1024, 1024, 8, 2, 128, is_causal=True, p_dropout=0.1, seed=78, offset=9), seed=13)


def test_static_scratch_backward_rejects_what_does_not_fit():
    b, s, h, d = 1, 4096, 32, 128                    # 4*1*32*4096*(1+128) = 64.5 MiB > 32 MiB
    assert 4 * b * h * s * (1 + d) > api.STATIC_SCRATCH_BYTES
    x = torch.zeros((b, s, h, d), dtype=torch.bfloat16, device="cuda")
    lse = torch.zeros((b, h, s), dtype=torch.float32, device="cuda")
    grads = (torch.full_like(x, 3.0), torch.full_like(x, 3.0), torch.full_like(x, 3.0))
    with pytest.raises(api.FlashAttnError, match="static array holds"):
        api.flash_attn_bwd_static_scratch(x, x, x, x, x, lse, grads=grads)
    torch.cuda.synchronize()
    assert all(bool((g == 3.0).all()) for g in grads)      # nothing was launched
