"""Ring context parallelism for sequences too long for one GPU (BASELINE.json config 5).

The only exchange step in the project.  The sequence is split over the N ranks of one NVLink box;
K/V chunks travel round a ring (torch.distributed P2P: NCCL over NVLink on GPUs, gloo in the CPU
tests) while each rank runs the ORDINARY flash_attn_fwd / flash_attn_bwd entry points
(/root/reference/flash_attn.h:10-36, :68-99) on (query slice, visiting K/V chunk) pairs.  The
exchange of step t+1 is issued before the kernels of step t, so it overlaps them.  Per-step results
are combined by the HBM-bound kernels of csrc/fa_cp.cu.

Causal attention uses zig-zag sharding: the sequence is cut into 2N chunks and rank r holds chunks
r and 2N-1-r, which balances the triangular work.  With that layout every ring step is one of
three dense calls (c = chunk length, src = owner of the visiting K/V):
  src == r : causal,      q[0:2c] x kv[0:2c]
  src <  r : non-causal,  q[0:2c] x kv[0:c]      (the visitor's late chunk is invisible)
  src >  r : non-causal,  q[c:2c] x kv[0:2c]     (only the late query chunk sees the visitor)
Non-causal attention takes any contiguous sharding and calls q x kv non-causally every step.

Dropout: every step is an ordinary call with the caller's seed, the offset `step_offset(offset, rank, src, world)` and
batch_base = the batch entry, rows and columns counted inside the step's (query slice, K/V chunk).  The mask is a function
of (seed, offset, world) -- not the unsharded call's -- and the backward of a step repeats its forward's
(include/flash_attn_b200_ext.h; restated independently in tests/ring_ref.py).

Two engines run this schedule.  On CUDA tensors (the product): the library's own C entry points
`flash_attn_b200_ring_fwd` / `flash_attn_b200_ring_bwd` (csrc/fa_ring.cu: ncclSend / ncclRecv on a communication
stream one step ahead of the kernels, event-chained; one merge kernel at the end of the forward) -- this module
then only bootstraps the communicator (`RingComm`) and owns the workspace.  With an explicit `ops` engine (the
CPU tests pass an oracle-backed one under gloo): the Python loop below over `torch.distributed` P2P, which is the
same schedule step for step.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

import ctypes

from .shard import CudaOps

ENGINE = "C entry points flash_attn_b200_ring_fwd / _bwd (csrc/fa_ring.cu) over ncclSend / ncclRecv"


class RingComm:
    """The library's ring communicator for one process group: NCCL unique id from rank 0, handed round with
    torch.distributed (any backend), then `flash_attn_b200_ring_comm_create` on every rank (its device current)."""

    def __init__(self, group=None):
        from . import api
        self.api, self.lib = api, api.lib()
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        L = self.lib
        vp, i, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
        L.flash_attn_b200_ring_unique_id.argtypes = [vp]; L.flash_attn_b200_ring_unique_id.restype = ctypes.c_bool
        L.flash_attn_b200_ring_comm_create.argtypes = [vp, i, i, ctypes.POINTER(vp)]
        L.flash_attn_b200_ring_comm_create.restype = ctypes.c_bool
        L.flash_attn_b200_ring_comm_destroy.argtypes = [vp]; L.flash_attn_b200_ring_comm_destroy.restype = None
        for name in ("fwd", "bwd"):
            f = getattr(L, f"flash_attn_b200_ring_{name}_workspace_bytes")
            f.argtypes = [i] * 6; f.restype = sz
        L.flash_attn_b200_ring_fwd.argtypes = [vp] * 7 + [sz] + [i] * 5 + [ctypes.c_float, ctypes.c_bool, ctypes.c_bool, vp]
        L.flash_attn_b200_ring_fwd.restype = ctypes.c_bool
        L.flash_attn_b200_ring_bwd.argtypes = [vp] * 11 + [sz] + [i] * 5 + [ctypes.c_float, ctypes.c_bool, ctypes.c_bool, vp]
        L.flash_attn_b200_ring_bwd.restype = ctypes.c_bool
        drop = [ctypes.c_float, ctypes.c_uint64, ctypes.c_uint64]           # p_dropout, seed, offset
        L.flash_attn_b200_ring_fwd_dropout.argtypes = L.flash_attn_b200_ring_fwd.argtypes + drop
        L.flash_attn_b200_ring_fwd_dropout.restype = ctypes.c_bool
        L.flash_attn_b200_ring_bwd_dropout.argtypes = L.flash_attn_b200_ring_bwd.argtypes + drop
        L.flash_attn_b200_ring_bwd_dropout.restype = ctypes.c_bool
        ident = [None]
        if self.world > 1:
            if self.rank == 0:
                buf = ctypes.create_string_buffer(128)
                if not L.flash_attn_b200_ring_unique_id(buf):
                    raise api.FlashAttnError(api.last_error())
                ident = [buf.raw]
            src = dist.get_global_rank(group, 0) if group is not None else 0
            dist.broadcast_object_list(ident, src=src, group=group)
        self.handle = vp()
        if not L.flash_attn_b200_ring_comm_create(ident[0], self.world, self.rank, ctypes.byref(self.handle)):
            raise api.FlashAttnError(api.last_error())
        self._ws = {}

    def workspace(self, kind, b, s, h, hk, d, device):
        n = getattr(self.lib, f"flash_attn_b200_ring_{kind}_workspace_bytes")(self.world, b, s, h, hk, d)
        ws = self._ws.get(kind)
        if ws is None or ws.numel() < n or ws.device != device:
            ws = self._ws[kind] = torch.empty(max(n, 256), dtype=torch.uint8, device=device)
        return ws

    def close(self):
        if self.handle:
            self.lib.flash_attn_b200_ring_comm_destroy(self.handle)
            self.handle = None


_COMMS = {}


def _comm_for(group) -> RingComm:
    key = id(group) if group is not None else None
    if key not in _COMMS:
        _COMMS[key] = RingComm(group)
    return _COMMS[key]


def close_comms() -> None:
    """Destroys the cached communicators (call before destroy_process_group)."""
    for c in _COMMS.values():
        c.close()
    _COMMS.clear()


def step_offset(offset: int, rank: int, src: int, world: int) -> int:
    """Philox offset of the ring step (rank, src): the pair index sits in bits 48.., i.e. in the Philox key
    (csrc/fa_ring.cu `step_offset`)."""
    return (int(offset) + ((rank * world + src) << 48)) & 0xFFFFFFFFFFFFFFFF


def _c_fwd(q, k, v, causal, scale, group, p_dropout=0.0, seed=0, offset=0):
    c = _comm_for(group)
    b, s, h, d = q.shape
    hk = k.shape[2]
    assert q.is_contiguous() and k.is_contiguous() and v.is_contiguous()
    out = torch.empty_like(q)
    lse = torch.empty((b, h, s), dtype=torch.float32, device=q.device)
    ws = c.workspace("fwd", b, s, h, hk, d, q.device)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    args = (c.handle, P(q), P(k), P(v), P(out), P(lse), P(ws), ws.numel(), b, s, h, hk, d, scale, causal,
            q.dtype == torch.bfloat16, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    if p_dropout > 0:
        ok = c.lib.flash_attn_b200_ring_fwd_dropout(*args, p_dropout, seed, offset)
    else:
        ok = c.lib.flash_attn_b200_ring_fwd(*args)
    if not ok:
        raise c.api.FlashAttnError("flash_attn_b200_ring_fwd failed: " + c.api.last_error())
    return out, lse


def _c_bwd(dout, q, k, v, out, lse, causal, scale, group, p_dropout=0.0, seed=0, offset=0):
    c = _comm_for(group)
    b, s, h, d = q.shape
    hk = k.shape[2]
    for t in (dout, q, k, v, out, lse):
        assert t.is_contiguous()
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    ws = c.workspace("bwd", b, s, h, hk, d, q.device)
    P = lambda t: ctypes.c_void_p(t.data_ptr())
    args = (c.handle, P(dout), P(q), P(k), P(v), P(out), P(lse), P(dq), P(dk), P(dv), P(ws), ws.numel(), b, s, h, hk, d,
            scale, causal, q.dtype == torch.bfloat16, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    if p_dropout > 0:
        ok = c.lib.flash_attn_b200_ring_bwd_dropout(*args, p_dropout, seed, offset)
    else:
        ok = c.lib.flash_attn_b200_ring_bwd(*args)
    if not ok:
        raise c.api.FlashAttnError("flash_attn_b200_ring_bwd failed: " + c.api.last_error())
    return dq, dk, dv


# ---- zig-zag layout helpers (used by callers, the bench and the tests) -------------------------

def zigzag_shard(x: torch.Tensor, rank: int, world_size: int, dim: int = 1) -> torch.Tensor:
    """Rank `rank`'s zig-zag shard of a full-sequence tensor: chunks rank and 2N-1-rank along `dim`."""
    chunks = x.chunk(2 * world_size, dim=dim)
    if len(chunks) != 2 * world_size or any(c.shape[dim] != chunks[0].shape[dim] for c in chunks):
        raise ValueError("sequence length must be a multiple of 2 * world_size")
    return torch.cat([chunks[rank], chunks[2 * world_size - 1 - rank]], dim=dim).contiguous()


def zigzag_unshard(shards: List[torch.Tensor], dim: int = 1) -> torch.Tensor:
    """Inverse of zigzag_shard over the list of all ranks' shards."""
    n = len(shards)
    chunks = [None] * (2 * n)
    for r, s in enumerate(shards):
        a, b = s.chunk(2, dim=dim)
        chunks[r], chunks[2 * n - 1 - r] = a, b
    return torch.cat(chunks, dim=dim)


def _exchange(send: torch.Tensor, recv: torch.Tensor, group) -> list:
    """send -> next rank, recv <- previous rank; returns the requests to wait on."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    nxt = dist.get_global_rank(group, (rank + 1) % world) if group is not None else (rank + 1) % world
    prv = dist.get_global_rank(group, (rank - 1) % world) if group is not None else (rank - 1) % world
    ops = [dist.P2POp(dist.isend, send, nxt, group), dist.P2POp(dist.irecv, recv, prv, group)]
    if rank % 2:                      # pairwise-ordered for backends that match P2P ops in issue order
        ops.reverse()
    return dist.batch_isend_irecv(ops)


def _wait(reqs) -> None:
    for r in reqs:
        r.wait()


def _steps(causal: bool, rank: int, src: int, s_local: int):
    """(q rows, kv rows, causal flag) of one ring step; row ranges are (start, stop) in local rows."""
    if not causal:
        return (0, s_local), (0, s_local), False
    c = s_local // 2
    if src == rank:
        return (0, s_local), (0, s_local), True
    if src < rank:
        return (0, s_local), (0, c), False
    return (c, s_local), (0, s_local), False


def ring_attn_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, *, causal: bool,
                  scale: Optional[float] = None, group=None, ops=None, p_dropout: float = 0.0, seed: int = 0,
                  offset: int = 0):
    """Forward over the ring.  q [b, s_local, h, d], k/v [b, s_local, hk, d]: this rank's shard
    (zig-zag when causal).  Returns (out [b, s_local, h, d], lse fp32 [b, h, s_local]) equal to the
    rank's rows of full-sequence attention."""
    b, s, h, d = q.shape
    if causal and s % 2:
        raise ValueError("causal ring attention needs an even local length (zig-zag chunks)")
    scale = float(d) ** -0.5 if scale is None else scale
    if ops is None and q.is_cuda:
        return _c_fwd(q, k, v, causal, scale, group, p_dropout, seed, offset)
    ops = ops if ops is not None else CudaOps()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    o_acc = torch.zeros((b, s, h, d), dtype=torch.float32, device=q.device)
    lse_acc = torch.full((b, h, s), float("inf"), dtype=torch.float32, device=q.device)
    kv_cur = torch.stack([k, v]).contiguous()
    kv_nxt = torch.empty_like(kv_cur)
    for t in range(world):
        reqs = _exchange(kv_cur, kv_nxt, group) if t + 1 < world else []
        src = (rank - t) % world
        (q0, q1), (k0, k1), step_causal = _steps(causal, rank, src, s)
        for bi in range(b):
            o_blk, lse_blk = ops.fwd(q[bi:bi + 1, q0:q1], kv_cur[0, bi:bi + 1, k0:k1], kv_cur[1, bi:bi + 1, k0:k1],
                                     causal=step_causal, scale=scale, p_dropout=p_dropout, seed=seed,
                                     offset=step_offset(offset, rank, src, world), batch_base=bi)
            ops.merge(o_acc[bi, q0:q1], lse_acc[bi, :, q0:q1], o_blk[0], lse_blk[0])
        _wait(reqs)
        kv_cur, kv_nxt = kv_nxt, kv_cur
    out = torch.empty_like(q)
    ops.store(out, o_acc)
    return out, lse_acc


def ring_attn_bwd(dout: torch.Tensor, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, out: torch.Tensor,
                  lse: torch.Tensor, *, causal: bool, scale: Optional[float] = None, group=None, ops=None,
                  p_dropout: float = 0.0, seed: int = 0, offset: int = 0):
    """Backward over the ring from the forward's (out, lse).  The dK/dV partial sums travel with
    their K/V chunk (fp32) and arrive home after the last hop.  Returns (dq, dk, dv)."""
    b, s, h, d = q.shape
    scale = float(d) ** -0.5 if scale is None else scale
    if ops is None and q.is_cuda:
        return _c_bwd(dout, q, k, v, out, lse, causal, scale, group, p_dropout, seed, offset)
    ops = ops if ops is not None else CudaOps()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dq_acc = torch.zeros((b, s, h, d), dtype=torch.float32, device=q.device)
    kv_cur = torch.stack([k, v]).contiguous()
    kv_nxt = torch.empty_like(kv_cur)
    dkv_cur = torch.zeros(kv_cur.shape, dtype=torch.float32, device=q.device)
    dkv_nxt = torch.empty_like(dkv_cur)
    dkv_reqs: list = []
    for t in range(world):
        kv_reqs = _exchange(kv_cur, kv_nxt, group) if t + 1 < world else []
        src = (rank - t) % world
        (q0, q1), (k0, k1), step_causal = _steps(causal, rank, src, s)
        blocks = []
        for bi in range(b):
            lse_b = lse[bi:bi + 1, :, q0:q1].contiguous()
            blocks.append(ops.bwd(dout[bi:bi + 1, q0:q1], q[bi:bi + 1, q0:q1], kv_cur[0, bi:bi + 1, k0:k1],
                                  kv_cur[1, bi:bi + 1, k0:k1], out[bi:bi + 1, q0:q1], lse_b,
                                  causal=step_causal, scale=scale, p_dropout=p_dropout, seed=seed,
                                  offset=step_offset(offset, rank, src, world), batch_base=bi))
        if t > 0:                     # partial sums for the chunk held now arrive from the previous rank
            _wait(dkv_reqs)
            dkv_cur, dkv_nxt = dkv_nxt, dkv_cur
        for bi, (dq_b, dk_b, dv_b) in enumerate(blocks):
            ops.accumulate(dq_acc[bi, q0:q1], dq_b[0])
            ops.accumulate(dkv_cur[0, bi, k0:k1], dk_b[0])
            ops.accumulate(dkv_cur[1, bi, k0:k1], dv_b[0])
        if world > 1:                 # pass them on; after the last step this hop is the way home
            dkv_reqs = _exchange(dkv_cur, dkv_nxt, group)
        _wait(kv_reqs)
        kv_cur, kv_nxt = kv_nxt, kv_cur
    if world > 1:
        _wait(dkv_reqs)
        dkv_cur = dkv_nxt
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    ops.store(dq, dq_acc)
    ops.store(dk, dkv_cur[0].contiguous())
    ops.store(dv, dkv_cur[1].contiguous())
    return dq, dk, dv
