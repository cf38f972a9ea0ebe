"""Batch sharding of the four entry points across the GPUs of one box.

Every (batch, head) pair is an independent problem and the reference layouts put batch outermost
(/root/reference/flash_attn.h:10-12), so a rank's shard is a pointer offset and a smaller
`batch_size` (:17): there is NO data-path collective.  The only cross-rank state is the dropout
stream, which is indexed by the global batch (csrc/philox_spec.h); the `batch_base` argument of the
`flash_attn_b200_*_shard` entry points (include/flash_attn_b200_ext.h) keeps a shard's masks equal to the
unsharded call's.

`ops` is the compute engine: the C-ABI library by default (`CudaOps`); the CPU tests pass an
oracle-backed engine so that the host logic runs under gloo without a GPU.
"""
from __future__ import annotations

from typing import Optional, Tuple


def batch_shard(global_batch: int, world_size: int, rank: int) -> Tuple[int, int]:
    """Contiguous split of `global_batch` over `world_size` ranks: (first batch, count) of `rank`.
    The first `global_batch % world_size` ranks hold one extra batch; a rank may hold none."""
    if world_size <= 0 or not 0 <= rank < world_size or global_batch < 0:
        raise ValueError("batch_shard: bad arguments")
    base, rem = divmod(global_batch, world_size)
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


class CudaOps:
    """The product engine: every call goes through libflash_attn.so (no fallback)."""

    def __init__(self):
        from . import api
        self.api = api

    def fwd(self, q, k, v, *, causal, scale, p_dropout=0.0, seed=0, offset=0, batch_base=0):
        out, lse, _, _ = self.api.flash_attn_fwd(q, k, v, p_dropout=p_dropout, softmax_scale=scale,
                                                 is_causal=causal, seed=seed, offset=offset, batch_base=batch_base)
        return out, lse

    def bwd(self, dout, q, k, v, out, lse, *, causal, scale, p_dropout=0.0, seed=0, offset=0, batch_base=0):
        return self.api.flash_attn_bwd(dout, q, k, v, out, lse, p_dropout=p_dropout, softmax_scale=scale,
                                       is_causal=causal, seed=seed, offset=offset, batch_base=batch_base)

    # ring context-parallel combine steps (csrc/fa_cp.cu)
    def merge(self, o_acc, lse_acc, o_blk, lse_blk):
        self.api.cp_merge(o_acc, lse_acc, o_blk, lse_blk)

    def accumulate(self, acc, blk):
        self.api.cp_accumulate(acc, blk)

    def store(self, dst, src):
        self.api.cp_store(dst, src)


class BatchShardedAttention:
    """One rank's view of a batch-sharded attention layer.

    q/k/v passed to fwd/bwd are this rank's batches [first, first + count) of the global tensors, in
    the reference's dense layout.  Results equal the corresponding slices of the unsharded call,
    dropout included.
    """

    def __init__(self, global_batch: int, rank: int, world_size: int, ops=None):
        self.first, self.count = batch_shard(global_batch, world_size, rank)
        self.ops = ops if ops is not None else CudaOps()

    def local_slice(self, x_global):
        """The rank's slice of a [global_batch, ...] tensor (a view: no copy)."""
        return x_global[self.first:self.first + self.count]

    def fwd(self, q, k, v, *, causal: bool, scale: Optional[float] = None, p_dropout: float = 0.0,
            seed: int = 0, offset: int = 0):
        if q.shape[0] != self.count:
            raise ValueError(f"expected {self.count} local batches, got {q.shape[0]}")
        if self.count == 0:
            return q.new_empty(q.shape), q.new_empty((0, q.shape[2], q.shape[1]))
        scale = float(q.shape[-1]) ** -0.5 if scale is None else scale
        return self.ops.fwd(q, k, v, causal=causal, scale=scale, p_dropout=p_dropout, seed=seed,
                            offset=offset, batch_base=self.first)

    def bwd(self, dout, q, k, v, out, lse, *, causal: bool, scale: Optional[float] = None,
            p_dropout: float = 0.0, seed: int = 0, offset: int = 0):
        if q.shape[0] != self.count:
            raise ValueError(f"expected {self.count} local batches, got {q.shape[0]}")
        if self.count == 0:
            return q.new_empty(q.shape), k.new_empty(k.shape), v.new_empty(v.shape)
        scale = float(q.shape[-1]) ** -0.5 if scale is None else scale
        return self.ops.bwd(dout, q, k, v, out, lse, causal=causal, scale=scale, p_dropout=p_dropout,
                            seed=seed, offset=offset, batch_base=self.first)
