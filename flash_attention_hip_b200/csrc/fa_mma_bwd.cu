// fa_mma_bwd.cu -- all-shapes backward (reference flash_attn.h:68-134) on warp-level tensor-core MMAs.
//
// Serves what the tcgen05 backward (fa_tc_bwd.cu) cannot: head sizes 136..256 (dK and dV alone would
// fill tensor memory), attn_mask together with packed inputs or num_splits == 1.  Two kernels, no
// atomics, so dQ is deterministic by construction:
//   * mma_bwd_dq_kernel   one CTA = 64 query rows of one (batch, head); loops over 64-key tiles;
//                         dQ (64 x D, fp32 fragments) stays in registers.
//   * mma_bwd_dkdv_kernel one CTA = 64 keys of one (batch, kv head); loops over the query heads of the
//                         GQA group and their 64-row tiles; dK and dV stay in registers.
// Both recompute S = Q K^T and dP = dO V^T for their (query tile, key tile) pairs, turn them into
// P (dropout applied, csrc/philox_spec.h: one Philox call per 16 columns) and dS in shared memory,
// and feed those to the accumulating GEMMs.  fp32 accumulation, P and dS rounded to the input dtype
// as in the tcgen05 kernels.  nvcuda::wmma 16x16x16 fragments (HMMA on sm_100a).
#include <mma.h>

#include "fa_common.cuh"

namespace fa {
namespace {

using namespace nvcuda;

constexpr int kB = 64;            // query rows / keys per tile
constexpr int kThreads = 256;     // 8 warps
constexpr int kLdS = 68;          // fp32 score tiles: row stride in floats
constexpr int kLdP = 72;          // 16-bit P / dS tiles: row stride in elements

template <int kD>
struct Smem {
  static constexpr int kLd = kD + 8;                        // 16-bit operand tiles: row stride in elements
  static constexpr size_t kTile = (size_t)kB * kLd * 2;     // bytes of one 64 x kD operand tile
  static constexpr size_t kQ = 0, kK = kTile, kV = 2 * kTile, kG = 3 * kTile;
  static constexpr size_t kS = 4 * kTile, kDP = kS + (size_t)kB * kLdS * 4;
  static constexpr size_t kP = kDP + (size_t)kB * kLdS * 4, kDS = kP + (size_t)kB * kLdP * 2;
  static constexpr size_t kBytes = kDS + (size_t)kB * kLdP * 2;
  static_assert(2 * kTile >= (size_t)kB * kD * 4, "the fp32 output staging aliases two operand tiles");
  static_assert(kBytes <= 227 * 1024, "shared memory");
};

// rows [row0, row0 + 64) of a [tokens, heads, d] tensor -> smem tile, zero past `rows` and past d
template <typename T, int kD>
__device__ __forceinline__ void load_tile(T *dst, const T *src, int row0, int rows, size_t row_stride, int d) {
  constexpr int kLd = Smem<kD>::kLd, kVec = kD / 8;
  for (int idx = threadIdx.x; idx < kB * kVec; idx += kThreads) {
    const int r = idx / kVec, c = (idx % kVec) * 8;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (row0 + r < rows && c < d) v = *reinterpret_cast<const uint4 *>(src + (size_t)(row0 + r) * row_stride + c);
    *reinterpret_cast<uint4 *>(dst + r * kLd + c) = v;
  }
}

// S = Q K^T and dP = dO V^T of one 64 x 64 pair -> fp32 tiles in shared memory (rows = queries)
template <typename T, int kD>
__device__ __forceinline__ void scores(const T *sQ, const T *sK, const T *sV, const T *sG, float *sS, float *sDP) {
  constexpr int kLd = Smem<kD>::kLd;
  const int warp = threadIdx.x >> 5;
  const int fr = warp >> 1, fc0 = (warp & 1) * 2;
  wmma::fragment<wmma::accumulator, 16, 16, 16, float> s[2], dp[2];
#pragma unroll
  for (int c = 0; c < 2; ++c) { wmma::fill_fragment(s[c], 0.f); wmma::fill_fragment(dp[c], 0.f); }
#pragma unroll 4
  for (int k = 0; k < kD / 16; ++k) {
    wmma::fragment<wmma::matrix_a, 16, 16, 16, T, wmma::row_major> a, g;
    wmma::load_matrix_sync(a, sQ + fr * 16 * kLd + k * 16, kLd);
    wmma::load_matrix_sync(g, sG + fr * 16 * kLd + k * 16, kLd);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      wmma::fragment<wmma::matrix_b, 16, 16, 16, T, wmma::col_major> bk, bv;
      wmma::load_matrix_sync(bk, sK + (fc0 + c) * 16 * kLd + k * 16, kLd);
      wmma::load_matrix_sync(bv, sV + (fc0 + c) * 16 * kLd + k * 16, kLd);
      wmma::mma_sync(s[c], a, bk, s[c]);
      wmma::mma_sync(dp[c], g, bv, dp[c]);
    }
  }
#pragma unroll
  for (int c = 0; c < 2; ++c) {
    wmma::store_matrix_sync(sS + fr * 16 * kLdS + (fc0 + c) * 16, s[c], kLdS, wmma::mem_row_major);
    wmma::store_matrix_sync(sDP + fr * 16 * kLdS + (fc0 + c) * 16, dp[c], kLdS, wmma::mem_row_major);
  }
}

// P (dropout applied and rescaled: the dV operand) and dS = scale * P o (dropout(dP) - D) as 16-bit tiles.
// Thread t owns row t / 4, columns 16 (t % 4) .. +15: exactly one Philox call (csrc/philox_spec.h).
template <typename T>
__device__ __forceinline__ void probs(const Params &p, const SeqInfo &si, int b, int h, int q0, int k0,
                                      const float *sS, const float *sDP, T *sP, T *sDS) {
  const int r = threadIdx.x >> 2, c0 = (threadIdx.x & 3) * 16;
  const int row = q0 + r;
  const bool row_ok = row < si.sq;
  const float L = row_ok ? p.lse[((size_t)b * p.h + h) * p.sq + row] : INFINITY;
  const float Dr = row_ok ? p.softmax_d[((size_t)b * p.h + h) * p.sq_rounded + row] : 0.f;
  const int last = p.causal ? min(si.sk - 1, row + si.sk - si.sq) : si.sk - 1;   // last visible key
  const bool drop = p.p_dropout > 0.f;
  fa_philox4_t rnd;
  rnd.v[0] = rnd.v[1] = rnd.v[2] = rnd.v[3] = 0;
  if (drop)
    rnd = fa_dropout_bytes16(p.seed, p.offset, (uint32_t)((p.batch_base + b) * p.h + h), (uint32_t)row,
                             (uint32_t)((k0 + c0) / FA_DROPOUT_COLS_PER_CALL));
  const T *bias = nullptr;
  if (p.mask && row_ok) {
    const int64_t bi = p.mask_b == 1 ? 0 : b, hi = p.mask_h == 1 ? 0 : h;
    bias = static_cast<const T *>(p.mask) + ((bi * p.mask_h + hi) * p.mask_sq + row) * p.mask_sk;
  }
#pragma unroll
  for (int e = 0; e < 16; ++e) {
    const int col = k0 + c0 + e;
    float pd = 0.f, ds = 0.f;
    if (row_ok && col <= last && L != INFINITY) {
      float s = sS[r * kLdS + c0 + e] * p.scale;
      if (bias) s += to_f32<T>(bias[col]);
      const float pn = __expf(s - L);
      float dp = sDP[r * kLdS + c0 + e];
      pd = pn;
      if (drop) {
        const bool keep = ((rnd.v[e >> 2] >> (8 * (e & 3))) & 0xffu) < p.drop_threshold;
        dp = keep ? dp * p.rp_dropout : 0.f;
        pd = keep ? pn * p.rp_dropout : 0.f;
      }
      ds = pn * (dp - Dr) * p.scale;
    }
    sP[r * kLdP + c0 + e] = from_f32<T>(pd);
    sDS[r * kLdP + c0 + e] = from_f32<T>(ds);
  }
}

// 64 x kD fp32 staging tile -> rows [row0, ...) of a [tokens, heads, d] output, bounded by `rows` and d
template <typename T, int kD>
__device__ __forceinline__ void store_tile(T *dst, const float *stage, int row0, int rows, size_t row_stride, int d) {
  constexpr int kVec = kD / 8;
  for (int idx = threadIdx.x; idx < kB * kVec; idx += kThreads) {
    const int r = idx / kVec, c = (idx % kVec) * 8;
    if (row0 + r >= rows || c >= d) continue;
    const float4 x = *reinterpret_cast<const float4 *>(stage + r * kD + c);
    const float4 y = *reinterpret_cast<const float4 *>(stage + r * kD + c + 4);
    T o[8] = {from_f32<T>(x.x), from_f32<T>(x.y), from_f32<T>(x.z), from_f32<T>(x.w),
              from_f32<T>(y.x), from_f32<T>(y.y), from_f32<T>(y.z), from_f32<T>(y.w)};
    *reinterpret_cast<uint4 *>(dst + (size_t)(row0 + r) * row_stride + c) = *reinterpret_cast<const uint4 *>(o);
  }
}

template <typename T, int kD>
__global__ void __launch_bounds__(kThreads, 1) mma_bwd_dq_kernel(const Params p) {
  using S = Smem<kD>;
  constexpr int kLd = S::kLd, kFrags = kD / 32;          // accumulator fragments per warp (half the columns)
  extern __shared__ __align__(128) uint8_t smem[];
  T *sQ = reinterpret_cast<T *>(smem + S::kQ), *sK = reinterpret_cast<T *>(smem + S::kK);
  T *sV = reinterpret_cast<T *>(smem + S::kV), *sG = reinterpret_cast<T *>(smem + S::kG);
  float *sS = reinterpret_cast<float *>(smem + S::kS), *sDP = reinterpret_cast<float *>(smem + S::kDP);
  T *sP = reinterpret_cast<T *>(smem + S::kP), *sDS = reinterpret_cast<T *>(smem + S::kDS);
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kB;
  const SeqInfo si = seq_info(p, b);
  if (q0 >= si.sq) return;
  const int hk = h / (p.h / p.hk), d = p.d;
  const size_t qs = (size_t)p.h * d, ks = (size_t)p.hk * d;
  const T *qb = static_cast<const T *>(p.q) + ((size_t)si.q0 * p.h + h) * d;
  const T *gb = static_cast<const T *>(p.dout) + ((size_t)si.q0 * p.h + h) * d;
  const T *kb = static_cast<const T *>(p.k) + ((size_t)si.k0 * p.hk + hk) * d;
  const T *vb = static_cast<const T *>(p.v) + ((size_t)si.k0 * p.hk + hk) * d;
  const int warp = threadIdx.x >> 5, fr = warp >> 1, col0 = (warp & 1) * (kD / 2);

  load_tile<T, kD>(sQ, qb, q0, si.sq, qs, d);
  load_tile<T, kD>(sG, gb, q0, si.sq, qs, d);
  wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[kFrags];
#pragma unroll
  for (int f = 0; f < kFrags; ++f) wmma::fill_fragment(acc[f], 0.f);

  int k_end = si.sk;                                     // keys any row of this tile can see
  if (p.causal) k_end = max(0, min(si.sk, min(q0 + kB, si.sq) + si.sk - si.sq));
  for (int k0 = 0; k0 < k_end; k0 += kB) {
    __syncthreads();                                     // the previous pair is done with sK / sV / sDS
    load_tile<T, kD>(sK, kb, k0, si.sk, ks, d);
    load_tile<T, kD>(sV, vb, k0, si.sk, ks, d);
    __syncthreads();
    scores<T, kD>(sQ, sK, sV, sG, sS, sDP);
    __syncthreads();
    probs<T>(p, si, b, h, q0, k0, sS, sDP, sP, sDS);
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < kB / 16; ++kk) {               // dQ += dS K
      wmma::fragment<wmma::matrix_a, 16, 16, 16, T, wmma::row_major> a;
      wmma::load_matrix_sync(a, sDS + fr * 16 * kLdP + kk * 16, kLdP);
#pragma unroll
      for (int f = 0; f < kFrags; ++f) {
        wmma::fragment<wmma::matrix_b, 16, 16, 16, T, wmma::row_major> bm;
        wmma::load_matrix_sync(bm, sK + kk * 16 * kLd + col0 + f * 16, kLd);
        wmma::mma_sync(acc[f], a, bm, acc[f]);
      }
    }
  }
  __syncthreads();
  float *stage = reinterpret_cast<float *>(smem + S::kK);   // sK and sV are free now
#pragma unroll
  for (int f = 0; f < kFrags; ++f)
    wmma::store_matrix_sync(stage + fr * 16 * kD + col0 + f * 16, acc[f], kD, wmma::mem_row_major);
  __syncthreads();
  store_tile<T, kD>(static_cast<T *>(p.dq) + ((size_t)si.q0 * p.h + h) * d, stage, q0, si.sq, qs, d);
}

template <typename T, int kD>
__global__ void __launch_bounds__(kThreads, 1) mma_bwd_dkdv_kernel(const Params p) {
  using S = Smem<kD>;
  constexpr int kLd = S::kLd, kFrags = kD / 32;
  extern __shared__ __align__(128) uint8_t smem[];
  T *sQ = reinterpret_cast<T *>(smem + S::kQ), *sK = reinterpret_cast<T *>(smem + S::kK);
  T *sV = reinterpret_cast<T *>(smem + S::kV), *sG = reinterpret_cast<T *>(smem + S::kG);
  float *sS = reinterpret_cast<float *>(smem + S::kS), *sDP = reinterpret_cast<float *>(smem + S::kDP);
  T *sP = reinterpret_cast<T *>(smem + S::kP), *sDS = reinterpret_cast<T *>(smem + S::kDS);
  const int b = blockIdx.z, hk = blockIdx.y, k0 = blockIdx.x * kB;
  const SeqInfo si = seq_info(p, b);
  if (k0 >= si.sk) return;
  const int group = p.h / p.hk, d = p.d;
  const size_t qs = (size_t)p.h * d, ks = (size_t)p.hk * d;
  const T *kb = static_cast<const T *>(p.k) + ((size_t)si.k0 * p.hk + hk) * d;
  const T *vb = static_cast<const T *>(p.v) + ((size_t)si.k0 * p.hk + hk) * d;
  const int warp = threadIdx.x >> 5, fr = warp >> 1, col0 = (warp & 1) * (kD / 2);

  load_tile<T, kD>(sK, kb, k0, si.sk, ks, d);
  load_tile<T, kD>(sV, vb, k0, si.sk, ks, d);
  wmma::fragment<wmma::accumulator, 16, 16, 16, float> dk[kFrags], dv[kFrags];
#pragma unroll
  for (int f = 0; f < kFrags; ++f) { wmma::fill_fragment(dk[f], 0.f); wmma::fill_fragment(dv[f], 0.f); }

  int q_begin = 0;                                       // first query tile that sees key k0
  if (p.causal) q_begin = max(0, k0 - (si.sk - si.sq)) / kB * kB;
  for (int h = hk * group; h < (hk + 1) * group; ++h) {
    const T *qb = static_cast<const T *>(p.q) + ((size_t)si.q0 * p.h + h) * d;
    const T *gb = static_cast<const T *>(p.dout) + ((size_t)si.q0 * p.h + h) * d;
    for (int q0 = q_begin; q0 < si.sq; q0 += kB) {
      __syncthreads();                                   // the previous pair is done with sQ / sG / sP / sDS
      load_tile<T, kD>(sQ, qb, q0, si.sq, qs, d);
      load_tile<T, kD>(sG, gb, q0, si.sq, qs, d);
      __syncthreads();
      scores<T, kD>(sQ, sK, sV, sG, sS, sDP);
      __syncthreads();
      probs<T>(p, si, b, h, q0, k0, sS, sDP, sP, sDS);
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < kB / 16; ++kk) {             // dV += P^T dO, dK += dS^T Q
        wmma::fragment<wmma::matrix_a, 16, 16, 16, T, wmma::col_major> ap, ad;
        wmma::load_matrix_sync(ap, sP + kk * 16 * kLdP + fr * 16, kLdP);
        wmma::load_matrix_sync(ad, sDS + kk * 16 * kLdP + fr * 16, kLdP);
#pragma unroll
        for (int f = 0; f < kFrags; ++f) {
          wmma::fragment<wmma::matrix_b, 16, 16, 16, T, wmma::row_major> bg, bq;
          wmma::load_matrix_sync(bg, sG + kk * 16 * kLd + col0 + f * 16, kLd);
          wmma::load_matrix_sync(bq, sQ + kk * 16 * kLd + col0 + f * 16, kLd);
          wmma::mma_sync(dv[f], ap, bg, dv[f]);
          wmma::mma_sync(dk[f], ad, bq, dk[f]);
        }
      }
    }
  }
  float *stage = reinterpret_cast<float *>(smem + S::kQ);   // sQ and sK (K is no longer needed either)
#pragma unroll
  for (int which = 0; which < 2; ++which) {
    __syncthreads();
#pragma unroll
    for (int f = 0; f < kFrags; ++f)
      wmma::store_matrix_sync(stage + fr * 16 * kD + col0 + f * 16, which ? dk[f] : dv[f], kD, wmma::mem_row_major);
    __syncthreads();
    T *dst = static_cast<T *>(which ? p.dk : p.dv) + ((size_t)si.k0 * p.hk + hk) * d;
    store_tile<T, kD>(dst, stage, k0, si.sk, ks, d);
  }
}

template <typename T, int kD>
const char *launch_t(const Params &p, cudaStream_t stream) {
  using S = Smem<kD>;
  auto kq = mma_bwd_dq_kernel<T, kD>;
  auto kk = mma_bwd_dkdv_kernel<T, kD>;
  int dev = 0, sm_count = 0;
  const char *de;
  if ((de = current_device(&dev, &sm_count))) return de;
  static std::atomic<uint64_t> opt_q{0}, opt_k{0};   // per instantiation, one bit per device
  if ((de = ensure_dynamic_smem(kq, (int)S::kBytes, dev, opt_q))) return de;
  if ((de = ensure_dynamic_smem(kk, (int)S::kBytes, dev, opt_k))) return de;
  if (p.sq > 0) kq<<<dim3((p.sq + kB - 1) / kB, p.h, p.b), kThreads, S::kBytes, stream>>>(p);
  if (p.sk > 0) kk<<<dim3((p.sk + kB - 1) / kB, p.hk, p.b), kThreads, S::kBytes, stream>>>(p);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

template <typename T>
const char *launch_d(const Params &p, cudaStream_t stream) {
  if (p.d <= 64) return launch_t<T, 64>(p, stream);
  if (p.d <= 128) return launch_t<T, 128>(p, stream);
  if (p.d <= 192) return launch_t<T, 192>(p, stream);
  return launch_t<T, 256>(p, stream);
}

}  // namespace

const char *launch_mma_bwd(const Params &p, cudaStream_t stream) {
  if (p.d <= 0 || p.d > 256 || p.d % 8) return "backward: head_size must be a multiple of 8 and <= 256";
  if (p.b == 0) return nullptr;
  return p.is_bf16 ? launch_d<__nv_bfloat16>(p, stream) : launch_d<__half>(p, stream);
}

}  // namespace fa
