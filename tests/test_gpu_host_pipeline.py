"""HostPipelinedAttention (host buffers, batch chunks on three streams) returns bit for bit what one
unchunked call on device tensors returns, dropout included (the chunks keep their global batch index
through the batch_base argument of the flash_attn_b200_*_shard entry points)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("b,chunks,hk,p_drop", [(6, 3, 4, 0.1), (5, 4, 2, 0.0), (2, 8, 4, 0.25)])
def test_pipelined_host_call_equals_one_shot(b, chunks, hk, p_drop):
    from flash_attention_hip_b200 import api
    from flash_attention_hip_b200.host_pipeline import HostPipelinedAttention
    s, h, d, base = 512, 4, 128, 7
    g = torch.Generator().manual_seed(11)
    mk = lambda heads: torch.randn((b, s, heads, d), generator=g).bfloat16().pin_memory()
    hq, hdo, hkk, hv = mk(h), mk(h), mk(hk), mk(hk)
    outs = [torch.empty_like(t).pin_memory() for t in (hq, hq, hkk, hkk)]
    pipe = HostPipelinedAttention(b, s, h, hk, d, chunks=chunks)
    for _ in range(2):                      # the second pass reuses staging buffers and scratch
        pipe.step(hq, hkk, hv, hdo, *outs, causal=True, p_dropout=p_drop, seed=3, offset=5, batch_base=base)
        torch.cuda.synchronize()
    assert pipe.launches > 0

    q, k, v, do = (t.cuda() for t in (hq, hkk, hv, hdo))
    kw = dict(p_dropout=p_drop, is_causal=True, seed=3, offset=5, batch_base=base)
    out, lse, _, _ = api.flash_attn_fwd(q, k, v, **kw)
    dq, dk, dv = api.flash_attn_bwd(do, q, k, v, out, lse, **kw)
    assert torch.equal(outs[0], out.cpu())
    # dQ is a sum of fp32 reduce-adds whose order is not fixed run to run: compare with a tolerance
    for got, ref in ((outs[1], dq), (outs[2], dk), (outs[3], dv)):
        ref = ref.float().cpu()
        assert float((got.float() - ref).abs().max()) <= 2e-2 * float(ref.abs().max())
