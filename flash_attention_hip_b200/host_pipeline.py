"""Forward + backward with HOST buffers, pipelined over batch chunks.

The four entry points take device pointers (/root/reference/flash_attn.h:10-16, :68-79).  A caller
whose tensors live in pinned host memory pays two PCIe transfers around them; since (batch, head)
units are independent, the batch can be cut into chunks and the three phases overlapped on three
streams: chunk c+1's host->device copy, chunk c's kernels, chunk c-1's device->host copy (PCIe is
full duplex, so the two copy directions overlap each other too).  Dropout stays identical to the
unchunked call through the `batch_base` argument of the `_shard` entry points.  This is what bench.py's `e2e` figure runs.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import api


class HostPipelinedAttention:
    """Device staging buffers + streams for one problem shape; reuse it across steps."""

    def __init__(self, b: int, s: int, h: int, hk: int, d: int, dtype=torch.bfloat16, chunks: int = 8,
                 device: Optional[torch.device] = None, fine_events: bool = True):
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.b, self.s, self.h, self.hk, self.d = b, s, h, hk, d
        self.bounds = [(b * c // chunks, b * (c + 1) // chunks) for c in range(chunks) if b * c // chunks < b * (c + 1) // chunks]
        mk = lambda heads: torch.empty((b, s, heads, d), dtype=dtype, device=self.device)
        self.q, self.do, self.out, self.dq = mk(h), mk(h), mk(h), mk(h)
        self.k, self.v, self.dk, self.dv = mk(hk), mk(hk), mk(hk), mk(hk)
        bc = max(i1 - i0 for i0, i1 in self.bounds)
        sq_r, d_r = api.round_up(s, 128), api.round_up(d, 32)
        # scratch is per chunk in flight: two sets, used alternately on the one compute stream
        self.scratch = [(torch.empty((bc, h, sq_r), dtype=torch.float32, device=self.device),
                         torch.empty((bc, sq_r, h, d_r), dtype=torch.float32, device=self.device)) for _ in range(2)]
        self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(device=self.device) for _ in range(3))
        self.launches = 0
        self.fine_events = fine_events      # False: one event per chunk and direction (kept for A/B timing)

    def step(self, hq, hk, hv, hdo, ho, hdq, hdk, hdv, *, causal: bool, p_dropout: float = 0.0, seed: int = 0,
             offset: int = 0, batch_base: int = 0, scale: Optional[float] = None) -> None:
        """hq..hdo: pinned host inputs [b, s, heads, d]; ho, hdq, hdk, hdv: pinned host outputs.
        Returns after everything is ENQUEUED; synchronise the device (or `self.s_out`) to read them."""
        cur = torch.cuda.current_stream(self.device)
        for st in (self.s_in, self.s_cmp, self.s_out):
            st.wait_stream(cur)
        for c, (i0, i1) in enumerate(self.bounds):
            # the forward needs q, k, v only and its result can leave while the backward runs: two events per direction
            # shorten the un-overlapped head (first chunk in) and tail (last chunk out) of the pipeline by a quarter each
            with torch.cuda.stream(self.s_in):
                for dst, src in ((self.q, hq), (self.k, hk), (self.v, hv)):
                    dst[i0:i1].copy_(src[i0:i1], non_blocking=True)
                qkv_ready = self.s_in.record_event()
                self.do[i0:i1].copy_(hdo[i0:i1], non_blocking=True)
                do_ready = self.s_in.record_event()
            kw = dict(p_dropout=p_dropout, softmax_scale=scale, is_causal=causal, seed=seed, offset=offset,
                      stream=self.s_cmp, batch_base=batch_base + i0)
            sd, acc = self.scratch[c & 1]
            n = i1 - i0
            # torch's allocator ties lse / rng_state to the stream that is current when they are made
            with torch.cuda.stream(self.s_cmp):
                self.s_cmp.wait_event(qkv_ready if self.fine_events else do_ready)
                _, lse, _, _ = api.flash_attn_fwd(self.q[i0:i1], self.k[i0:i1], self.v[i0:i1], out=self.out[i0:i1], **kw)
                self.launches += api.last_launch_count()
                fwd_done = self.s_cmp.record_event()
                self.s_cmp.wait_event(do_ready)
                api.flash_attn_bwd(self.do[i0:i1], self.q[i0:i1], self.k[i0:i1], self.v[i0:i1], self.out[i0:i1], lse,
                                   scratch=(sd[:n], acc[:n]), grads=(self.dq[i0:i1], self.dk[i0:i1], self.dv[i0:i1]),
                                   **kw)
                self.launches += api.last_launch_count()
                bwd_done = self.s_cmp.record_event()
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(fwd_done if self.fine_events else bwd_done)
                ho[i0:i1].copy_(self.out[i0:i1], non_blocking=True)
                self.s_out.wait_event(bwd_done)
                for dst, src in ((hdq, self.dq), (hdk, self.dk), (hdv, self.dv)):
                    dst[i0:i1].copy_(src[i0:i1], non_blocking=True)
        cur.wait_stream(self.s_out)
