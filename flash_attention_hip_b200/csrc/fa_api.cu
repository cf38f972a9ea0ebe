// fa_api.cu -- the C ABI of include/flash_attn.h: validation, parameter block,
// kernel-family choice.  Each entry point replaces the reference prototype of
// the same name (reference flash_attn.h:10, :38, :68, :101, :136, :138).
//
// There is no CPU fallback: every path ends in a CUDA launch or in `false`.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/flash_attn.h"
#include "../../include/flash_attn_b200_ext.h"
#include "fa_common.cuh"

namespace {

thread_local std::string g_error;
thread_local int g_path_request = FA_B200_PATH_AUTO;
thread_local int g_last_path = 0;
thread_local int g_last_launches = 0;

bool fail(const char *fmt, ...) __attribute__((format(printf, 1, 2)));
bool fail(const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  flash_attn_set_error(buf);
  return false;
}

bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

__global__ void write_rng_state_kernel(uint64_t *dst, uint64_t seed, uint64_t offset) {
  dst[0] = seed;
  dst[1] = offset;
}

struct Shape {
  int b, sq, sk, sq_r, sk_r, h, hk, d, d_r;
};

// Checks shared by the four entry points.  Returns false after set_error.
bool fill_common(fa::Params &p, const char *fn, const Shape &s, float p_dropout, float softmax_scale,
                 bool is_causal, bool is_bf16, uint64_t seed, uint64_t offset,
                 const int32_t *cu_q, const int32_t *cu_k, bool varlen, const void *attn_mask,
                 const int64_t *mask_dims, int batch_base) {
  if (s.b < 0 || s.sq < 0 || s.sk < 0) return fail("%s: negative batch_size or seqlen", fn);
  if (s.h <= 0 || s.hk <= 0) return fail("%s: num_heads and num_heads_k must be positive", fn);
  if (s.h % s.hk) return fail("%s: num_heads (%d) must be a multiple of num_heads_k (%d)", fn, s.h, s.hk);
  if (s.d <= 0 || s.d > 256 || s.d % 8)
    return fail("%s: head_size must be a multiple of 8 in [8, 256], got %d", fn, s.d);
  if (!(p_dropout >= 0.f && p_dropout <= 1.f)) return fail("%s: p_dropout must be in [0, 1]", fn);
  if (!std::isfinite(softmax_scale)) return fail("%s: softmax_scale must be finite", fn);
  if (varlen && (!cu_q || !cu_k)) return fail("%s: cu_seqlens_q / cu_seqlens_k must not be NULL", fn);
  if (s.b > 65535) return fail("%s: batch_size > 65535 is not supported", fn);
  if (batch_base < 0) return fail("%s: batch_base must not be negative", fn);
  std::memset(&p, 0, sizeof(p));
  p.b = s.b; p.sq = s.sq; p.sk = s.sk; p.sq_rounded = s.sq_r; p.sk_rounded = s.sk_r;
  p.h = s.h; p.hk = s.hk; p.d = s.d; p.d_rounded = s.d_r;
  p.scale = softmax_scale;
  p.p_dropout = p_dropout;
  p.rp_dropout = fa_dropout_rescale(p_dropout);
  p.drop_threshold = fa_dropout_threshold(p_dropout);
  p.batch_base = batch_base;
  p.seed = seed; p.offset = offset;
  p.causal = is_causal; p.is_bf16 = is_bf16;
  p.cu_q = varlen ? cu_q : nullptr;
  p.cu_k = varlen ? cu_k : nullptr;
  if (attn_mask) {
    if (!mask_dims) return fail("%s: attn_mask given but mask_dims is NULL", fn);
    if ((mask_dims[0] != 1 && mask_dims[0] != s.b) || (mask_dims[1] != 1 && mask_dims[1] != s.h) ||
        mask_dims[2] != s.sq || mask_dims[3] != s.sk)
      return fail("%s: mask_dims must be {1|batch, 1|num_heads, seqlen_q, seqlen_k} (packed inputs: the max_seqlen pair)", fn);
    p.mask = attn_mask;
    p.mask_b = mask_dims[0]; p.mask_h = mask_dims[1]; p.mask_sq = mask_dims[2]; p.mask_sk = mask_dims[3];
  }
  return true;
}

bool fwd_impl(const char *fn, const void *q, const void *k, const void *v, const int32_t *cu_q,
              const int32_t *cu_k, bool varlen, void *rng_state, void *out, void *softmax_ptr,
              void *softmax_lse_ptr, const Shape &s, float p_dropout, float softmax_scale,
              bool is_causal, bool return_softmax, bool is_bf16, fa_stream_t stream, uint64_t seed,
              uint64_t offset, const void *attn_mask, const int64_t *mask_dims, int batch_base = 0) {
  fa::Params p;
  if (!fill_common(p, fn, s, p_dropout, softmax_scale, is_causal, is_bf16, seed, offset, cu_q, cu_k,
                   varlen, attn_mask, mask_dims, batch_base))
    return false;
  const bool empty = s.b == 0 || s.sq == 0;
  if (!empty && (!q || !k || !v || !out || !softmax_lse_ptr))
    return fail("%s: q, k, v, out and softmax_lse_ptr must not be NULL", fn);
  if (!aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(out))
    return fail("%s: q, k, v, out must be 16-byte aligned", fn);
  if (return_softmax) {
    if (!softmax_ptr) return fail("%s: return_softmax set but softmax_ptr is NULL", fn);
    if (s.sq_r < s.sq || s.sk_r < s.sk)
      return fail("%s: seqlen_q_rounded/seqlen_k_rounded smaller than the sequence lengths", fn);
  }
  p.q = q; p.k = k; p.v = v; p.out = out;
  p.lse = static_cast<float *>(softmax_lse_ptr);
  p.softmax_ret = return_softmax ? softmax_ptr : nullptr;

  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool tc_ok = !empty && fa::tc_fwd_supported(p);
  if (!empty && g_path_request == FA_B200_PATH_TCGEN05 && !tc_ok)
    return fail("%s: tcgen05 path requested but this problem is outside it", fn);
  // every check has passed: only now is caller memory written
  int launches = 0;
  if (rng_state) {
    write_rng_state_kernel<<<1, 1, 0, st>>>(static_cast<uint64_t *>(rng_state), seed, offset);
    ++launches;
  }
  if (empty) { g_last_launches = launches; return true; }
  const bool use_tc = tc_ok && g_path_request != FA_B200_PATH_GENERIC;
  const char *err = use_tc ? fa::launch_tc_fwd(p, st) : fa::launch_generic_fwd(p, st);
  ++launches;
  // return_softmax beside the tcgen05 forward: a second pass over Q K^T from the LSE just written (the all-shapes
  // kernel writes it itself)
  if (!err && use_tc && p.softmax_ret) { err = fa::launch_softmax_ret(p, st); ++launches; }
  if (err) return fail("%s: %s", fn, err);
  g_last_path = use_tc ? FA_B200_PATH_TCGEN05 : FA_B200_PATH_GENERIC;
  g_last_launches = launches;
  return true;
}

bool bwd_impl(const char *fn, const void *dout, const void *q, const void *k, const void *v,
              const void *out, const void *softmax_d, const void *softmax_lse, const int32_t *cu_q,
              const int32_t *cu_k, bool varlen, void *dq, void *dk, void *dv, void *dq_accum,
              const Shape &s, float p_dropout, float softmax_scale, bool is_causal, bool is_bf16,
              int num_splits, fa_stream_t stream, uint64_t seed, uint64_t offset,
              const void *attn_mask, const int64_t *mask_dims, int batch_base = 0) {
  fa::Params p;
  if (!fill_common(p, fn, s, p_dropout, softmax_scale, is_causal, is_bf16, seed, offset, cu_q, cu_k,
                   varlen, attn_mask, mask_dims, batch_base))
    return false;
  if (s.b == 0) { g_last_launches = 0; return true; }
  if (!dout || !q || !k || !v || !out || !softmax_d || !softmax_lse || !dq || !dk || !dv)
    return fail("%s: dout, q, k, v, out, softmax_d, softmax_lse, dq, dk, dv must not be NULL", fn);
  if (!aligned16(dout) || !aligned16(q) || !aligned16(k) || !aligned16(v) || !aligned16(out) ||
      !aligned16(dq) || !aligned16(dk) || !aligned16(dv))
    return fail("%s: tensors must be 16-byte aligned", fn);
  if (s.sq_r < s.sq) return fail("%s: seqlen_q_rounded smaller than seqlen_q", fn);
  p.dout = dout; p.q = q; p.k = k; p.v = v; p.out = const_cast<void *>(out);
  p.softmax_d = const_cast<float *>(static_cast<const float *>(softmax_d));
  p.lse = const_cast<float *>(static_cast<const float *>(softmax_lse));
  p.dq = dq; p.dk = dk; p.dv = dv;
  p.dq_accum = static_cast<float *>(dq_accum);
  p.deterministic = num_splits == 1;

  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool tc_ok = fa::tc_bwd_supported(p);
  if (g_path_request == FA_B200_PATH_TCGEN05 && !tc_ok)
    return fail("%s: tcgen05 path requested but this problem is outside it", fn);
  const bool use_tc = tc_ok && g_path_request != FA_B200_PATH_GENERIC;
  // softmax_d; the tcgen05 backward's dq_accum is zeroed by the same kernel
  // (the two-launch backward -- num_splits == 1, attn_mask -- writes dq without the scratch: nothing to zero)
  const bool zero_acc = use_tc && fa::tc_bwd_uses_scratch(p);
  const char *err = fa::launch_bwd_preprocess(p, st, zero_acc ? p.dq_accum : nullptr, zero_acc ? fa::tc_bwd_scratch_floats(p) : 0);
  if (!err) err = use_tc ? fa::launch_tc_bwd(p, st) : fa::launch_generic_bwd(p, st);
  if (err) return fail("%s: %s", fn, err);
  g_last_path = use_tc ? FA_B200_PATH_TCGEN05 : FA_B200_PATH_GENERIC;
  g_last_launches = 3;
  return true;
}

}  // namespace

extern "C" {

// Replaces reference flash_attn.h:10-36.
bool flash_attn_fwd(const void *const q, const void *const k, const void *const v,
                    void *const rng_state, void *const out, void *const softmax_ptr,
                    void *const softmax_lse_ptr, const int batch_size, const int seqlen_q,
                    const int seqlen_k, const int seqlen_q_rounded, const int seqlen_k_rounded,
                    const int num_heads, const int num_heads_k, const int head_size,
                    const int head_size_rounded, const float p_dropout, const float softmax_scale,
                    const float softmax_unscale, const bool is_causal, const bool return_softmax,
                    const bool is_bf16, fa_stream_t stream, uint64_t seed, uint64_t offset,
                    const void *const attn_mask, const int64_t *const mask_dims) {
  (void)softmax_unscale;  // the bias is added after scaling q.k, so 1/scale is never needed
  try {
    const Shape s{batch_size, seqlen_q, seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return fwd_impl("flash_attn_fwd", q, k, v, nullptr, nullptr, false, rng_state, out, softmax_ptr,
                    softmax_lse_ptr, s, p_dropout, softmax_scale, is_causal, return_softmax, is_bf16,
                    stream, seed, offset, attn_mask, mask_dims);
  } catch (...) {
    return fail("flash_attn_fwd: unexpected exception");
  }
}

// Replaces reference flash_attn.h:38-66.
bool flash_attn_varlen_fwd(const void *const q, const void *const k, const void *const v,
                           const int32_t *const cu_seqlens_q, const int32_t *const cu_seqlens_k,
                           void *const rng_state, void *const out, void *const softmax_ptr,
                           void *const softmax_lse_ptr, const int batch_size,
                           const int max_seqlen_q, const int max_seqlen_k,
                           const int seqlen_q_rounded, const int seqlen_k_rounded,
                           const int num_heads, const int num_heads_k, const int head_size,
                           const int head_size_rounded, const float p_dropout,
                           const float softmax_scale, const float softmax_unscale,
                           const bool is_causal, const bool return_softmax, const bool is_bf16,
                           fa_stream_t stream, uint64_t seed, uint64_t offset,
                           const void *const attn_mask, const void *const mask_dims) {
  (void)softmax_unscale;
  try {
    const Shape s{batch_size, max_seqlen_q, max_seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return fwd_impl("flash_attn_varlen_fwd", q, k, v, cu_seqlens_q, cu_seqlens_k, true, rng_state,
                    out, softmax_ptr, softmax_lse_ptr, s, p_dropout, softmax_scale, is_causal,
                    return_softmax, is_bf16, stream, seed, offset, attn_mask,
                    static_cast<const int64_t *>(mask_dims));
  } catch (...) {
    return fail("flash_attn_varlen_fwd: unexpected exception");
  }
}

// Replaces reference flash_attn.h:68-99.
bool flash_attn_bwd(const void *const dout, const void *const q, const void *const k,
                    const void *const v, const void *const out, const void *const softmax_d,
                    const void *const softmax_lse, void *const rng_state, void *const dq,
                    void *const dk, void *const dv, void *const dq_accum, const int batch_size,
                    const int seqlen_q, const int seqlen_k, const int seqlen_q_rounded,
                    const int seqlen_k_rounded, const int num_heads, const int num_heads_k,
                    const int head_size, const int head_size_rounded, const float p_dropout,
                    const float softmax_scale, const float softmax_unscale, const bool is_causal,
                    const bool is_bf16, const int num_splits, fa_stream_t stream, uint64_t seed,
                    uint64_t offset, const void *const attn_mask, const int64_t *const mask_dims) {
  (void)softmax_unscale;
  (void)rng_state;  // seed/offset arrive by value (reference flash_attn.h:96-97)
  try {
    const Shape s{batch_size, seqlen_q, seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return bwd_impl("flash_attn_bwd", dout, q, k, v, out, softmax_d, softmax_lse, nullptr, nullptr,
                    false, dq, dk, dv, dq_accum, s, p_dropout, softmax_scale, is_causal, is_bf16,
                    num_splits, stream, seed, offset, attn_mask, mask_dims);
  } catch (...) {
    return fail("flash_attn_bwd: unexpected exception");
  }
}

// Replaces reference flash_attn.h:101-134.
bool flash_attn_varlen_bwd(const void *const dout, const void *const q, const void *const k,
                           const void *const v, const void *const out,
                           const void *const softmax_d, const void *const softmax_lse,
                           const int32_t *const cu_seqlens_q, const int32_t *const cu_seqlens_k,
                           void *const rng_state, void *const dq, void *const dk, void *const dv,
                           void *const dq_accum, const int batch_size, const int max_seqlen_q,
                           const int max_seqlen_k, const int seqlen_q_rounded,
                           const int seqlen_k_rounded, const int num_heads, const int num_heads_k,
                           const int head_size, const int head_size_rounded, const float p_dropout,
                           const float softmax_scale, const float softmax_unscale,
                           const bool is_causal, const bool is_bf16, const int num_splits,
                           fa_stream_t stream, uint64_t seed, uint64_t offset,
                           const void *attn_mask, const int64_t *const mask_dims) {
  (void)softmax_unscale;
  (void)rng_state;
  try {
    const Shape s{batch_size, max_seqlen_q, max_seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return bwd_impl("flash_attn_varlen_bwd", dout, q, k, v, out, softmax_d, softmax_lse,
                    cu_seqlens_q, cu_seqlens_k, true, dq, dk, dv, dq_accum, s, p_dropout,
                    softmax_scale, is_causal, is_bf16, num_splits, stream, seed, offset, attn_mask,
                    mask_dims);
  } catch (...) {
    return fail("flash_attn_varlen_bwd: unexpected exception");
  }
}

// Replaces reference flash_attn.h:136.  Thread-local.
void flash_attn_set_error(const char *msg) {
  try {
    g_error = msg ? msg : "";
  } catch (...) {
  }
}

// Replaces reference flash_attn.h:138.
const char *flash_attn_error(void) { return g_error.c_str(); }


// ---- batch-shard entry points (include/flash_attn_b200_ext.h): the reference calls plus an explicit batch_base --------
bool flash_attn_b200_fwd_shard(int batch_base, const void *q, const void *k, const void *v, void *rng_state, void *out,
                               void *softmax_ptr, void *softmax_lse_ptr, int batch_size, int seqlen_q, int seqlen_k,
                               int seqlen_q_rounded, int seqlen_k_rounded, int num_heads, int num_heads_k, int head_size,
                               int head_size_rounded, float p_dropout, float softmax_scale, float softmax_unscale,
                               bool is_causal, bool return_softmax, bool is_bf16, fa_stream_t stream, uint64_t seed,
                               uint64_t offset, const void *attn_mask, const int64_t *mask_dims) {
  (void)softmax_unscale;
  try {
    const Shape s{batch_size, seqlen_q, seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return fwd_impl("flash_attn_b200_fwd_shard", q, k, v, nullptr, nullptr, false, rng_state, out, softmax_ptr,
                    softmax_lse_ptr, s, p_dropout, softmax_scale, is_causal, return_softmax, is_bf16, stream, seed,
                    offset, attn_mask, mask_dims, batch_base);
  } catch (...) {
    return fail("flash_attn_b200_fwd_shard: unexpected exception");
  }
}

bool flash_attn_b200_varlen_fwd_shard(int batch_base, const void *q, const void *k, const void *v,
                                      const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, void *rng_state,
                                      void *out, void *softmax_ptr, void *softmax_lse_ptr, int batch_size,
                                      int max_seqlen_q, int max_seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
                                      int num_heads, int num_heads_k, int head_size, int head_size_rounded,
                                      float p_dropout, float softmax_scale, float softmax_unscale, bool is_causal,
                                      bool return_softmax, bool is_bf16, fa_stream_t stream, uint64_t seed,
                                      uint64_t offset, const void *attn_mask, const void *mask_dims) {
  (void)softmax_unscale;
  try {
    const Shape s{batch_size, max_seqlen_q, max_seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return fwd_impl("flash_attn_b200_varlen_fwd_shard", q, k, v, cu_seqlens_q, cu_seqlens_k, true, rng_state, out,
                    softmax_ptr, softmax_lse_ptr, s, p_dropout, softmax_scale, is_causal, return_softmax, is_bf16,
                    stream, seed, offset, attn_mask, static_cast<const int64_t *>(mask_dims), batch_base);
  } catch (...) {
    return fail("flash_attn_b200_varlen_fwd_shard: unexpected exception");
  }
}

bool flash_attn_b200_bwd_shard(int batch_base, const void *dout, const void *q, const void *k, const void *v,
                               const void *out, const void *softmax_d, const void *softmax_lse, void *rng_state,
                               void *dq, void *dk, void *dv, void *dq_accum, int batch_size, int seqlen_q, int seqlen_k,
                               int seqlen_q_rounded, int seqlen_k_rounded, int num_heads, int num_heads_k, int head_size,
                               int head_size_rounded, float p_dropout, float softmax_scale, float softmax_unscale,
                               bool is_causal, bool is_bf16, int num_splits, fa_stream_t stream, uint64_t seed,
                               uint64_t offset, const void *attn_mask, const int64_t *mask_dims) {
  (void)softmax_unscale;
  (void)rng_state;
  try {
    const Shape s{batch_size, seqlen_q, seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return bwd_impl("flash_attn_b200_bwd_shard", dout, q, k, v, out, softmax_d, softmax_lse, nullptr, nullptr, false,
                    dq, dk, dv, dq_accum, s, p_dropout, softmax_scale, is_causal, is_bf16, num_splits, stream, seed,
                    offset, attn_mask, mask_dims, batch_base);
  } catch (...) {
    return fail("flash_attn_b200_bwd_shard: unexpected exception");
  }
}

bool flash_attn_b200_varlen_bwd_shard(int batch_base, const void *dout, const void *q, const void *k, const void *v,
                                      const void *out, const void *softmax_d, const void *softmax_lse,
                                      const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, void *rng_state,
                                      void *dq, void *dk, void *dv, void *dq_accum, int batch_size, int max_seqlen_q,
                                      int max_seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded, int num_heads,
                                      int num_heads_k, int head_size, int head_size_rounded, float p_dropout,
                                      float softmax_scale, float softmax_unscale, bool is_causal, bool is_bf16,
                                      int num_splits, fa_stream_t stream, uint64_t seed, uint64_t offset,
                                      const void *attn_mask, const int64_t *mask_dims) {
  (void)softmax_unscale;
  (void)rng_state;
  try {
    const Shape s{batch_size, max_seqlen_q, max_seqlen_k, seqlen_q_rounded, seqlen_k_rounded,
                  num_heads, num_heads_k, head_size, head_size_rounded};
    return bwd_impl("flash_attn_b200_varlen_bwd_shard", dout, q, k, v, out, softmax_d, softmax_lse, cu_seqlens_q,
                    cu_seqlens_k, true, dq, dk, dv, dq_accum, s, p_dropout, softmax_scale, is_causal, is_bf16,
                    num_splits, stream, seed, offset, attn_mask, mask_dims, batch_base);
  } catch (...) {
    return fail("flash_attn_b200_varlen_bwd_shard: unexpected exception");
  }
}

// ---- backward without any allocation (include/flash_attn_b200_ext.h) ---------------------------------
// softmax_d and dq_accum, the two scratch tensors flash_attn_bwd takes from its caller, live in ONE file-scope
// device array of FA_B200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  In host
// code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is fetched once
// per device with cudaGetSymbolAddress (that device current) and handed to the kernels as a plain pointer.
static __device__ __align__(256) float g_static_scratch[FA_B200_STATIC_SCRATCH_BYTES / sizeof(float)];

static const char *static_scratch_base(float **base) {
  static std::atomic<float *> cache[fa::kMaxCachedDevices];   // zero-initialised; re-fetching is harmless
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return "cudaGetDevice failed";
  const bool cached = dev >= 0 && dev < fa::kMaxCachedDevices;
  float *ptr = cached ? cache[dev].load(std::memory_order_acquire) : nullptr;
  if (!ptr) {
    if (cudaGetSymbolAddress(reinterpret_cast<void **>(&ptr), g_static_scratch) != cudaSuccess || !ptr)
      return "cudaGetSymbolAddress(static scratch) failed";
    if (cached) cache[dev].store(ptr, std::memory_order_release);
  }
  *base = ptr;
  return nullptr;
}

bool flash_attn_b200_bwd_static_scratch(const void *dout, const void *q, const void *k, const void *v,
                                        const void *out, const void *softmax_lse, void *dq, void *dk, void *dv,
                                        int batch_size, int seqlen_q, int seqlen_k, int num_heads, int num_heads_k,
                                        int head_size, float p_dropout, float softmax_scale, bool is_causal,
                                        bool is_bf16, fa_stream_t stream, uint64_t seed, uint64_t offset) {
  const char *fn = "flash_attn_b200_bwd_static_scratch";
  try {
    if (batch_size < 0 || seqlen_q < 0 || seqlen_k < 0 || num_heads <= 0 || head_size <= 0)
      return fail("%s: bad shape", fn);
    const long long sq_r = (static_cast<long long>(seqlen_q) + 127) / 128 * 128;
    const long long d_r = (static_cast<long long>(head_size) + 31) / 32 * 32;
    // softmax_d [b, h, sq_rounded] first (padded to a 256-byte boundary), then dq_accum [b, sq_rounded, h, d_rounded]
    const long long d_floats = (static_cast<long long>(batch_size) * num_heads * sq_r + 63) / 64 * 64;
    const long long acc_floats = static_cast<long long>(batch_size) * sq_r * num_heads * d_r;
    const long long need = (d_floats + acc_floats) * static_cast<long long>(sizeof(float));
    if (need > static_cast<long long>(FA_B200_STATIC_SCRATCH_BYTES))
      return fail("%s: this problem needs %lld bytes of scratch (softmax_d + dq_accum = 4*batch*heads*"
                  "seqlen_q_rounded*(1 + head_size_rounded)), the static array holds %lld: use flash_attn_bwd "
                  "with caller-owned scratch", fn, need, static_cast<long long>(FA_B200_STATIC_SCRATCH_BYTES));
    float *base = nullptr;
    if (const char *e = static_scratch_base(&base)) return fail("%s: %s", fn, e);
    const Shape s{batch_size, seqlen_q, seqlen_k, static_cast<int>(sq_r), (seqlen_k + 127) / 128 * 128,
                  num_heads, num_heads_k, head_size, static_cast<int>(d_r)};
    // num_splits = 0: dQ by reduce-adds in arrival order, the header's default
    return bwd_impl(fn, dout, q, k, v, out, base, softmax_lse, nullptr, nullptr, false, dq, dk, dv,
                    base + d_floats, s, p_dropout, softmax_scale, is_causal, is_bf16, /*num_splits=*/0, stream,
                    seed, offset, nullptr, nullptr);
  } catch (...) {
    return fail("%s: unexpected exception", fn);
  }
}

void flash_attn_b200_set_path(int path) { g_path_request = path; }
int flash_attn_b200_last_path(void) { return g_last_path; }
int flash_attn_b200_last_launch_count(void) { return g_last_launches; }
int flash_attn_b200_trace_read(int kernel, int role, unsigned long long *dst, int max_events) {
  return kernel == 0 ? fa::trace_read_fwd(dst, role, max_events) : fa::trace_read_bwd(dst, role, max_events);
}
// ---- ring context-parallel helpers (include/flash_attn_b200_ext.h; csrc/fa_cp.cu) ----------------
bool flash_attn_b200_cp_merge(void *o_acc, void *lse_acc, const void *o_blk, const void *lse_blk, int rows,
                              int num_heads, int head_size, int64_t lse_acc_stride, int64_t lse_blk_stride,
                              bool is_bf16, fa_stream_t stream) {
  if (!o_acc || !lse_acc || !o_blk || !lse_blk) return fail("flash_attn_b200_cp_merge: null pointer");
  if (rows < 0 || num_heads <= 0 || head_size <= 0) return fail("flash_attn_b200_cp_merge: bad shape");
  const char *e = fa::launch_cp_merge(o_acc, lse_acc, o_blk, lse_blk, rows, num_heads, head_size, lse_acc_stride,
                                      lse_blk_stride, is_bf16, reinterpret_cast<cudaStream_t>(stream));
  return e ? fail("flash_attn_b200_cp_merge: %s", e) : true;
}
bool flash_attn_b200_cp_accumulate(void *acc_f32, const void *blk, int64_t n, bool is_bf16, fa_stream_t stream) {
  if (!acc_f32 || !blk || n < 0) return fail("flash_attn_b200_cp_accumulate: bad argument");
  const char *e = fa::launch_cp_accumulate(acc_f32, blk, n, is_bf16, reinterpret_cast<cudaStream_t>(stream));
  return e ? fail("flash_attn_b200_cp_accumulate: %s", e) : true;
}
bool flash_attn_b200_cp_store(void *dst, const void *src_f32, int64_t n, bool is_bf16, fa_stream_t stream) {
  if (!dst || !src_f32 || n < 0) return fail("flash_attn_b200_cp_store: bad argument");
  const char *e = fa::launch_cp_store(dst, src_f32, n, is_bf16, reinterpret_cast<cudaStream_t>(stream));
  return e ? fail("flash_attn_b200_cp_store: %s", e) : true;
}

const char *flash_attn_b200_version(void) { return "flash_attn_b200 0.1 (sm_100a)"; }

}  // extern "C"
