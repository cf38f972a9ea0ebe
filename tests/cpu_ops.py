"""Oracle-backed compute engine for the CPU (gloo) tests of the multi-GPU host logic.

TEST INFRASTRUCTURE: it stands in for flash_attention_hip_b200.shard.CudaOps so that the ring /
batch-shard schedules can run without a GPU.  The product never imports this.
"""
import numpy as np
import torch

from oracle import oracle


def _bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


def _from_bits(a: np.ndarray, dtype) -> torch.Tensor:
    return torch.from_numpy(a.view(np.int16).copy()).view(dtype)


class OracleOps:
    def _problem(self, q, k, causal, scale, p_dropout, seed, offset, batch_base):
        b, sq, h, d = q.shape
        _, sk, hk, _ = k.shape
        return oracle.Problem(batch=b, seqlen_q=sq, seqlen_k=sk, num_heads=h, num_heads_k=hk, head_size=d,
                              is_causal=causal, is_bf16=q.dtype == torch.bfloat16, p_dropout=p_dropout,
                              softmax_scale=scale, seed=seed, offset=offset, batch_base=batch_base)

    def fwd(self, q, k, v, *, causal, scale, p_dropout=0.0, seed=0, offset=0, batch_base=0):
        pb = self._problem(q, k, causal, scale, p_dropout, seed, offset, batch_base)
        r = oracle.forward(pb, _bits(q).reshape(pb.q_shape()), _bits(k).reshape(pb.k_shape()),
                           _bits(v).reshape(pb.k_shape()))
        return _from_bits(r["out"], q.dtype).reshape(q.shape), torch.from_numpy(r["lse"])

    def bwd(self, dout, q, k, v, out, lse, *, causal, scale, p_dropout=0.0, seed=0, offset=0, batch_base=0):
        pb = self._problem(q, k, causal, scale, p_dropout, seed, offset, batch_base)
        r = oracle.backward(pb, _bits(dout).reshape(pb.q_shape()), _bits(q).reshape(pb.q_shape()),
                            _bits(k).reshape(pb.k_shape()), _bits(v).reshape(pb.k_shape()),
                            _bits(out).reshape(pb.q_shape()), lse.contiguous().numpy())
        return (_from_bits(r["dq"], q.dtype).reshape(q.shape), _from_bits(r["dk"], q.dtype).reshape(k.shape),
                _from_bits(r["dv"], q.dtype).reshape(v.shape))

    # the combine steps, restating csrc/fa_cp.cu
    def merge(self, o_acc, lse_acc, o_blk, lse_blk):
        la, lb = lse_acc.clone(), lse_blk
        empty_b, empty_a = torch.isinf(lb) & (lb > 0), torch.isinf(la) & (la > 0)
        la_f = torch.where(empty_a, torch.full_like(la, -float("inf")), la)
        lb_f = torch.where(empty_b, torch.full_like(lb, -float("inf")), lb)
        ln = torch.logaddexp(la_f, lb_f)
        wa = torch.where(empty_a, torch.zeros_like(la), torch.exp(la_f - ln))
        wb = torch.where(empty_b, torch.zeros_like(lb), torch.exp(lb_f - ln))
        both_empty = empty_a & empty_b
        wa = torch.where(both_empty, torch.ones_like(wa), wa)
        wb = torch.where(both_empty, torch.zeros_like(wb), wb)
        ln = torch.where(both_empty, torch.full_like(ln, float("inf")), ln)
        # [h, rows] -> [rows, h, 1]
        o_acc.copy_(wa.t().unsqueeze(-1) * o_acc + wb.t().unsqueeze(-1) * o_blk.float())
        lse_acc.copy_(ln)

    def accumulate(self, acc, blk):
        acc.add_(blk.float())

    def store(self, dst, src):
        dst.copy_(src.to(dst.dtype))
