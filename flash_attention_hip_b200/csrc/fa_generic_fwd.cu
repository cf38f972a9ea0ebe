// fa_generic_fwd.cu -- all-shapes forward kernel (plain SIMT, fp32 math).
//
// A second implementation of the forward of the flash_attn.h ABI (reference
// flash_attn.h:10-66) for the tests (flash_attn_b200_set_path(GENERIC)) and for
// a non-finite softmax_scale: any head size, attn_mask bias, return_softmax.  One warp owns one query row; lanes own keys
// for the scores and output columns for P.V, so every global read is either a
// full contiguous row per lane (K) or a coalesced row segment per warp (V).
#include "fa_common.cuh"

namespace fa {

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kMaxD = 256;

// dot(qs[0..d), row[0..d)), row read as 16-byte vectors (d % 8 == 0).
template <typename T>
__device__ __forceinline__ float dot_row(const float *qs, const T *row, int d) {
  float acc = 0.f;
  const uint4 *r4 = reinterpret_cast<const uint4 *>(row);
  for (int c8 = 0; c8 < d / 8; ++c8) {
    const uint4 raw = r4[c8];
    const T *e = reinterpret_cast<const T *>(&raw);
#pragma unroll
    for (int t = 0; t < 8; ++t) acc = fmaf(qs[c8 * 8 + t], to_f32<T>(e[t]), acc);
  }
  return acc;
}

template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
generic_fwd_kernel(const Params p) {
  __shared__ float qs_all[kWarpsPerBlock][kMaxD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y;
  const int row = blockIdx.x * kWarpsPerBlock + warp;
  const SeqInfo si = seq_info(p, b);
  if (row >= si.sq) return;  // whole warp leaves together
  const int d = p.d;
  const int hk = h / (p.h / p.hk);
  const T *q = static_cast<const T *>(p.q) + ((size_t)(si.q0 + row) * p.h + h) * d;
  const T *kbase = static_cast<const T *>(p.k) + ((size_t)si.k0 * p.hk + hk) * d;
  const T *vbase = static_cast<const T *>(p.v) + ((size_t)si.k0 * p.hk + hk) * d;
  const size_t kv_stride = (size_t)p.hk * d;
  float *qs = qs_all[warp];
  for (int c = lane; c < d; c += 32) qs[c] = to_f32<T>(q[c]);
  __syncwarp();

  // bias row (attn_mask, reference flash_attn.h:35-36)
  const T *bias = nullptr;
  if (p.mask) {
    const int64_t bi = p.mask_b == 1 ? 0 : b, hi = p.mask_h == 1 ? 0 : h;
    bias = static_cast<const T *>(p.mask) + ((bi * p.mask_h + hi) * p.mask_sq + row) * p.mask_sk;
  }

  // visible keys: [0, nvis)
  int nvis = si.sk;
  if (p.causal) nvis = min(si.sk, row + si.sk - si.sq + 1);
  if (nvis < 0) nvis = 0;

  const uint32_t bh = (uint32_t)((p.batch_base + b) * p.h + h);
  const bool drop = p.p_dropout > 0.f;
  float m = -INFINITY, l = 0.f;
  float o[kMaxD / 32];
#pragma unroll
  for (int t = 0; t < kMaxD / 32; ++t) o[t] = 0.f;

  for (int kb = 0; kb < nvis; kb += 32) {
    const int key = kb + lane;
    float s = -INFINITY;
    if (key < nvis) {
      s = dot_row<T>(qs, kbase + (size_t)key * kv_stride, d) * p.scale;
      if (bias) s += to_f32<T>(bias[key]);
    }
    const float m_new = fmaxf(m, warp_max(s));
    // m_new can stay -inf only if every score so far is -inf (bias of -inf)
    const float m_use = m_new == -INFINITY ? 0.f : m_new;
    const float pr = key < nvis ? __expf(s - m_use) : 0.f;
    const float corr = m == -INFINITY ? 0.f : __expf(m - m_use);
    l = l * corr + warp_sum(pr);
    m = m_new;
    float pd = pr;
    if (drop && key < nvis)
      pd = fa_dropout_keep(p.seed, p.offset, bh, (uint32_t)row, (uint32_t)key, p.drop_threshold) ? pr : 0.f;
#pragma unroll
    for (int t = 0; t < kMaxD / 32; ++t) o[t] *= corr;
    const int nk = min(32, nvis - kb);
    for (int kk = 0; kk < nk; ++kk) {
      const float pk = __shfl_sync(0xffffffffu, pd, kk);
      const T *vr = vbase + (size_t)(kb + kk) * kv_stride;
#pragma unroll
      for (int t = 0; t < kMaxD / 32; ++t) {
        const int c = lane + 32 * t;
        if (c < d) o[t] = fmaf(pk, to_f32<T>(vr[c]), o[t]);
      }
    }
  }

  const float inv = l > 0.f ? p.rp_dropout / l : 0.f;
  T *out = static_cast<T *>(p.out) + ((size_t)(si.q0 + row) * p.h + h) * d;
#pragma unroll
  for (int t = 0; t < kMaxD / 32; ++t) {
    const int c = lane + 32 * t;
    if (c < d) out[c] = from_f32<T>(o[t] * inv);
  }
  const float lse = l > 0.f ? m + __logf(l) : INFINITY;
  if (lane == 0) p.lse[((size_t)b * p.h + h) * p.sq + row] = lse;

  // Optional S_dmask output (reference flash_attn.h:15,:30): normalised
  // probabilities, negated where dropped.  Second pass, debug-grade speed.
  if (p.softmax_ret) {
    T *sr = static_cast<T *>(p.softmax_ret) +
            (((size_t)b * p.h + h) * p.sq_rounded + row) * (size_t)p.sk_rounded;
    for (int key = lane; key < si.sk; key += 32) {
      float pn = 0.f;
      if (key < nvis && l > 0.f) {
        float s = dot_row<T>(qs, kbase + (size_t)key * kv_stride, d) * p.scale;
        if (bias) s += to_f32<T>(bias[key]);
        pn = __expf(s - lse);
        if (drop && !fa_dropout_keep(p.seed, p.offset, bh, (uint32_t)row, (uint32_t)key, p.drop_threshold))
          pn = -pn;
      }
      sr[key] = from_f32<T>(pn);
    }
  }
}

}  // namespace

const char *launch_generic_fwd(const Params &p, cudaStream_t stream) {
  if (p.d > kMaxD || p.d % 8) return "generic forward: head_size must be a multiple of 8 and <= 256";
  if (p.b == 0 || p.sq == 0) return nullptr;
  dim3 grid((p.sq + kWarpsPerBlock - 1) / kWarpsPerBlock, p.h, p.b);
  if (p.is_bf16)
    generic_fwd_kernel<__nv_bfloat16><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p);
  else
    generic_fwd_kernel<__half><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace fa
