"""Host-side mirror of the flash_attn.h operator interface (ctypes over the C ABI).

Every function here ends in one call through libflash_attn.so's exported
symbols -- the same six names the reference header declares
(/root/reference/flash_attn.h:10-138).  torch is used only for device memory
and streams.  There is no CPU path: a missing library or a failed call raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Tuple

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# developer builds: FA_B200_TRACE=1 -> in-kernel event timelines (build.py --trace);
# FA_B200_VARIANT=x -> a tuning variant (build.py --variant=x -D...)
_VARIANT = "trace" if os.environ.get("FA_B200_TRACE") == "1" else os.environ.get("FA_B200_VARIANT", "")
LIB_PATH = os.path.join(_HERE, "lib", "libflash_attn" + ("_" + _VARIANT if _VARIANT else "") + ".so")

PATH_AUTO, PATH_GENERIC, PATH_TCGEN05 = 0, 1, 2

_lib = None

_vp, _i, _f, _b, _u64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_bool, ctypes.c_uint64

# argument lists in the reference's order (reference flash_attn.h:10-36, :38-66, :68-99, :101-134)
_FWD = [_vp] * 7 + [_i] * 9 + [_f] * 3 + [_b] * 3 + [_vp, _u64, _u64, _vp, _vp]
_VFWD = [_vp] * 9 + [_i] * 9 + [_f] * 3 + [_b] * 3 + [_vp, _u64, _u64, _vp, _vp]
_BWD = [_vp] * 12 + [_i] * 9 + [_f] * 3 + [_b] * 2 + [_i, _vp, _u64, _u64, _vp, _vp]
_VBWD = [_vp] * 14 + [_i] * 9 + [_f] * 3 + [_b] * 2 + [_i, _vp, _u64, _u64, _vp, _vp]
assert (len(_FWD), len(_VFWD), len(_BWD), len(_VBWD)) == (27, 29, 32, 34)


class FlashAttnError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Loads the C-ABI library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FlashAttnError(
                f"{LIB_PATH} is missing: run `python -m flash_attention_hip_b200.build` "
                "(there is no fallback path)")
        l = ctypes.CDLL(LIB_PATH)
        for name, args in (("flash_attn_fwd", _FWD), ("flash_attn_varlen_fwd", _VFWD),
                           ("flash_attn_bwd", _BWD), ("flash_attn_varlen_bwd", _VBWD)):
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = ctypes.c_bool
            # batch-shard variants (include/flash_attn_b200_ext.h): the same call with a leading batch_base
            fn = getattr(l, name.replace("flash_attn_", "flash_attn_b200_") + "_shard")
            fn.argtypes = [_i] + args
            fn.restype = ctypes.c_bool
        l.flash_attn_error.restype = ctypes.c_char_p
        l.flash_attn_set_error.argtypes = [ctypes.c_char_p]
        l.flash_attn_b200_version.restype = ctypes.c_char_p
        l.flash_attn_b200_set_path.argtypes = [_i]
        _lib = l
    return _lib


def last_error() -> str:
    return lib().flash_attn_error().decode()


def set_path(path: int) -> None:
    lib().flash_attn_b200_set_path(path)


def last_path() -> int:
    return int(lib().flash_attn_b200_last_path())


def last_launch_count() -> int:
    return int(lib().flash_attn_b200_last_launch_count())


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _check(ok: bool, what: str):
    if not ok:
        raise FlashAttnError(f"{what} failed: {last_error()}")


def _stream_ptr(stream: Optional[torch.cuda.Stream]):
    s = stream if stream is not None else torch.cuda.current_stream()
    return ctypes.c_void_p(s.cuda_stream)


def _call(name: str, batch_base: int, *args) -> bool:
    """The reference entry point `name`, or its `_shard` variant when the caller is a batch shard."""
    if batch_base:
        return getattr(lib(), name.replace("flash_attn_", "flash_attn_b200_") + "_shard")(int(batch_base), *args)
    return getattr(lib(), name)(*args)


def _mask_args(attn_mask: Optional[torch.Tensor]):
    if attn_mask is None:
        return None, None, None
    assert attn_mask.dim() == 4 and attn_mask.is_contiguous()
    dims = (ctypes.c_int64 * 4)(*attn_mask.shape)
    return _ptr(attn_mask), ctypes.cast(dims, ctypes.c_void_p), dims


def flash_attn_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, *, p_dropout: float = 0.0,
                   softmax_scale: Optional[float] = None, is_causal: bool = False,
                   return_softmax: bool = False, seed: int = 0, offset: int = 0,
                   attn_mask: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
                   lse: Optional[torch.Tensor] = None, stream: Optional[torch.cuda.Stream] = None,
                   batch_base: int = 0):
    """Dense forward.  q [b, sq, h, d]; k, v [b, sk, hk, d]; fp16 or bf16, contiguous.
    Returns (out, softmax_lse [b, h, sq] fp32, softmax or None, rng_state [2] int64)."""
    assert q.is_cuda and q.is_contiguous() and k.is_contiguous() and v.is_contiguous()
    assert q.dtype in (torch.float16, torch.bfloat16) and k.dtype == q.dtype and v.dtype == q.dtype
    b, sq, h, d = q.shape
    _, sk, hk, _ = k.shape
    scale = float(d) ** -0.5 if softmax_scale is None else float(softmax_scale)
    sq_r, sk_r, d_r = round_up(sq, 128), round_up(sk, 128), round_up(d, 32)
    if out is None:
        out = torch.empty_like(q)
    if lse is None:
        lse = torch.empty((b, h, sq), dtype=torch.float32, device=q.device)
    sm = torch.zeros((b, h, sq_r, sk_r), dtype=q.dtype, device=q.device) if return_softmax else None
    rng = torch.empty(2, dtype=torch.int64, device=q.device)
    mptr, mdims, _keep = _mask_args(attn_mask)
    ok = _call("flash_attn_fwd", batch_base, _ptr(q), _ptr(k), _ptr(v), _ptr(rng), _ptr(out), _ptr(sm), _ptr(lse),
               b, sq, sk, sq_r, sk_r, h, hk, d, d_r, p_dropout, scale,
               1.0 / scale if scale != 0 else 0.0, is_causal, return_softmax,
               q.dtype == torch.bfloat16, _stream_ptr(stream), seed, offset, mptr, mdims)
    _check(ok, "flash_attn_fwd")
    return out, lse, sm, rng


def flash_attn_varlen_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                          cu_seqlens_q: torch.Tensor, cu_seqlens_k: torch.Tensor,
                          max_seqlen_q: int, max_seqlen_k: int, *, p_dropout: float = 0.0,
                          softmax_scale: Optional[float] = None, is_causal: bool = False,
                          return_softmax: bool = False, seed: int = 0, offset: int = 0,
                          out: Optional[torch.Tensor] = None, lse: Optional[torch.Tensor] = None,
                          stream: Optional[torch.cuda.Stream] = None, attn_mask: Optional[torch.Tensor] = None,
                          batch_base: int = 0):
    """Packed forward.  q [total_q, h, d]; k, v [total_k, hk, d]; cu_seqlens int32 [b+1] on device.
    Returns (out, softmax_lse [b, h, max_seqlen_q], softmax or None, rng_state)."""
    assert q.is_cuda and q.is_contiguous() and k.is_contiguous() and v.is_contiguous()
    assert cu_seqlens_q.dtype == torch.int32 and cu_seqlens_k.dtype == torch.int32
    _, h, d = q.shape
    _, hk, _ = k.shape
    b = cu_seqlens_q.numel() - 1
    scale = float(d) ** -0.5 if softmax_scale is None else float(softmax_scale)
    sq_r, sk_r, d_r = round_up(max_seqlen_q, 128), round_up(max_seqlen_k, 128), round_up(d, 32)
    if out is None:
        out = torch.empty_like(q)
    if lse is None:
        lse = torch.full((b, h, max_seqlen_q), float("inf"), dtype=torch.float32, device=q.device)
    sm = torch.zeros((b, h, sq_r, sk_r), dtype=q.dtype, device=q.device) if return_softmax else None
    rng = torch.empty(2, dtype=torch.int64, device=q.device)
    mptr, mdims, _keep = _mask_args(attn_mask)
    ok = _call("flash_attn_varlen_fwd", batch_base, _ptr(q), _ptr(k), _ptr(v), _ptr(cu_seqlens_q), _ptr(cu_seqlens_k),
               _ptr(rng), _ptr(out), _ptr(sm), _ptr(lse), b, max_seqlen_q,
               max_seqlen_k, sq_r, sk_r, h, hk, d, d_r, p_dropout, scale,
               1.0 / scale if scale != 0 else 0.0, is_causal, return_softmax,
               q.dtype == torch.bfloat16, _stream_ptr(stream), seed, offset, mptr, mdims)
    _check(ok, "flash_attn_varlen_fwd")
    return out, lse, sm, rng


def _bwd_scratch(b, sq_r, h, d_r, device):
    softmax_d = torch.empty((b, h, sq_r), dtype=torch.float32, device=device)
    dq_accum = torch.empty((b, sq_r, h, d_r), dtype=torch.float32, device=device)
    return softmax_d, dq_accum


def flash_attn_bwd(dout, q, k, v, out, softmax_lse, *, p_dropout: float = 0.0,
                   softmax_scale: Optional[float] = None, is_causal: bool = False,
                   deterministic: bool = False, seed: int = 0, offset: int = 0,
                   attn_mask: Optional[torch.Tensor] = None, scratch=None,
                   grads: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = None,
                   stream: Optional[torch.cuda.Stream] = None, batch_base: int = 0):
    """Dense backward; returns (dq, dk, dv).  seed/offset/batch_base must be the forward's."""
    for t in (dout, q, k, v, out):
        assert t.is_cuda and t.is_contiguous()
    b, sq, h, d = q.shape
    _, sk, hk, _ = k.shape
    scale = float(d) ** -0.5 if softmax_scale is None else float(softmax_scale)
    sq_r, sk_r, d_r = round_up(sq, 128), round_up(sk, 128), round_up(d, 32)
    softmax_d, dq_accum = scratch if scratch is not None else _bwd_scratch(b, sq_r, h, d_r, q.device)
    dq, dk, dv = grads if grads is not None else (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    mptr, mdims, _keep = _mask_args(attn_mask)
    ok = _call("flash_attn_bwd", batch_base, _ptr(dout), _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(softmax_d),
               _ptr(softmax_lse), None, _ptr(dq), _ptr(dk), _ptr(dv), _ptr(dq_accum),
               b, sq, sk, sq_r, sk_r, h, hk, d, d_r, p_dropout, scale,
               1.0 / scale if scale != 0 else 0.0, is_causal,
               q.dtype == torch.bfloat16, 1 if deterministic else 0,
               _stream_ptr(stream), seed, offset, mptr, mdims)
    _check(ok, "flash_attn_bwd")
    return dq, dk, dv


def flash_attn_varlen_bwd(dout, q, k, v, out, softmax_lse, cu_seqlens_q, cu_seqlens_k,
                          max_seqlen_q: int, max_seqlen_k: int, *, p_dropout: float = 0.0,
                          softmax_scale: Optional[float] = None, is_causal: bool = False,
                          deterministic: bool = False, seed: int = 0, offset: int = 0, scratch=None,
                          grads: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = None,
                          stream: Optional[torch.cuda.Stream] = None, attn_mask: Optional[torch.Tensor] = None,
                          batch_base: int = 0):
    """Packed backward; returns (dq, dk, dv)."""
    for t in (dout, q, k, v, out):
        assert t.is_cuda and t.is_contiguous()
    _, h, d = q.shape
    _, hk, _ = k.shape
    b = cu_seqlens_q.numel() - 1
    scale = float(d) ** -0.5 if softmax_scale is None else float(softmax_scale)
    sq_r, sk_r, d_r = round_up(max_seqlen_q, 128), round_up(max_seqlen_k, 128), round_up(d, 32)
    softmax_d, dq_accum = scratch if scratch is not None else _bwd_scratch(b, sq_r, h, d_r, q.device)
    dq, dk, dv = grads if grads is not None else (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    mptr, mdims, _keep = _mask_args(attn_mask)
    ok = _call("flash_attn_varlen_bwd", batch_base, _ptr(dout), _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(softmax_d),
               _ptr(softmax_lse), _ptr(cu_seqlens_q), _ptr(cu_seqlens_k), None,
               _ptr(dq), _ptr(dk), _ptr(dv), _ptr(dq_accum), b, max_seqlen_q,
               max_seqlen_k, sq_r, sk_r, h, hk, d, d_r, p_dropout, scale,
               1.0 / scale if scale != 0 else 0.0, is_causal,
               q.dtype == torch.bfloat16, 1 if deterministic else 0,
               _stream_ptr(stream), seed, offset, mptr, mdims)
    _check(ok, "flash_attn_varlen_bwd")
    return dq, dk, dv


# ---- extensions beside the reference ABI (include/flash_attn_b200_ext.h) ---------------------

_i64 = ctypes.c_int64


def _ext(name, argtypes):
    fn = getattr(lib(), name)
    fn.argtypes = argtypes
    fn.restype = ctypes.c_bool
    return fn


def cp_merge(o_acc: torch.Tensor, lse_acc: torch.Tensor, o_blk: torch.Tensor, lse_blk: torch.Tensor,
             stream: Optional[torch.cuda.Stream] = None) -> None:
    """Ring step merge.  o_acc fp32 [rows, h, d] (contiguous view), lse_acc fp32 [h, rows] view whose
    last dim is contiguous (row stride = lse_acc.stride(0)); o_blk 16-bit [rows, h, d]; lse_blk fp32
    [h, rows].  lse == +inf marks an empty row."""
    rows, h, d = o_acc.shape
    assert o_acc.dtype == torch.float32 and o_acc.is_contiguous() and o_blk.is_contiguous()
    assert o_blk.shape == o_acc.shape and lse_acc.shape == (h, rows) and lse_blk.shape == (h, rows)
    assert lse_acc.stride(1) == 1 and lse_blk.stride(1) == 1
    fn = _ext("flash_attn_b200_cp_merge", [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _i64, _b, _vp])
    ok = fn(_ptr(o_acc), _ptr(lse_acc), _ptr(o_blk), _ptr(lse_blk), rows, h, d, lse_acc.stride(0),
            lse_blk.stride(0), o_blk.dtype == torch.bfloat16, _stream_ptr(stream))
    _check(ok, "flash_attn_b200_cp_merge")


def cp_accumulate(acc: torch.Tensor, blk: torch.Tensor, stream: Optional[torch.cuda.Stream] = None) -> None:
    """acc (fp32, contiguous) += blk (fp16/bf16, contiguous, same element count)."""
    assert acc.dtype == torch.float32 and acc.is_contiguous() and blk.is_contiguous()
    assert acc.numel() == blk.numel()
    fn = _ext("flash_attn_b200_cp_accumulate", [_vp, _vp, _i64, _b, _vp])
    _check(fn(_ptr(acc), _ptr(blk), acc.numel(), blk.dtype == torch.bfloat16, _stream_ptr(stream)),
           "flash_attn_b200_cp_accumulate")


def cp_store(dst: torch.Tensor, src: torch.Tensor, stream: Optional[torch.cuda.Stream] = None) -> None:
    """dst (fp16/bf16, contiguous) = src (fp32, contiguous)."""
    assert src.dtype == torch.float32 and src.is_contiguous() and dst.is_contiguous()
    assert dst.numel() == src.numel()
    fn = _ext("flash_attn_b200_cp_store", [_vp, _vp, _i64, _b, _vp])
    _check(fn(_ptr(dst), _ptr(src), src.numel(), dst.dtype == torch.bfloat16, _stream_ptr(stream)),
           "flash_attn_b200_cp_store")


STATIC_SCRATCH_BYTES = 32 << 20   # FA_B200_STATIC_SCRATCH_BYTES (include/flash_attn_b200_ext.h)


def flash_attn_bwd_static_scratch(dout, q, k, v, out, softmax_lse, *, p_dropout: float = 0.0,
                                  softmax_scale: Optional[float] = None, is_causal: bool = False,
                                  seed: int = 0, offset: int = 0,
                                  grads: Optional[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = None,
                                  stream: Optional[torch.cuda.Stream] = None):
    """Dense backward through `flash_attn_b200_bwd_static_scratch`: the library's own static device array is the
    scratch, the call allocates nothing (pass `grads` to keep torch from allocating the outputs either).
    Raises FlashAttnError when the problem needs more than STATIC_SCRATCH_BYTES."""
    for t in (dout, q, k, v, out):
        assert t.is_cuda and t.is_contiguous()
    b, sq, h, d = q.shape
    _, sk, hk, _ = k.shape
    scale = float(d) ** -0.5 if softmax_scale is None else float(softmax_scale)
    dq, dk, dv = grads if grads is not None else (torch.empty_like(q), torch.empty_like(k), torch.empty_like(v))
    fn = _ext("flash_attn_b200_bwd_static_scratch",
              [_vp] * 9 + [_i] * 6 + [_f, _f, _b, _b, _vp, _u64, _u64])
    ok = fn(_ptr(dout), _ptr(q), _ptr(k), _ptr(v), _ptr(out), _ptr(softmax_lse), _ptr(dq), _ptr(dk), _ptr(dv),
            b, sq, sk, h, hk, d, p_dropout, scale, is_causal, q.dtype == torch.bfloat16, _stream_ptr(stream),
            seed, offset)
    _check(ok, "flash_attn_b200_bwd_static_scratch")
    return dq, dk, dv
