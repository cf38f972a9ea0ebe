// fa_generic_bwd.cu -- all-shapes backward kernels (plain SIMT, fp32 math).
//
// Serves flash_attn_bwd / flash_attn_varlen_bwd (reference flash_attn.h:68-134)
// for whatever the tcgen05 backward does not specialise.  Three launches:
//   preprocess : softmax_d[b,h,i] = sum_c dout*out            (shared with tc path)
//   dq         : one warp per query row, lanes over keys
//   dk/dv      : one warp per key row of one kv head, lanes over query rows,
//                looping over the query heads of the GQA group
// P is recomputed from the stored LSE.  No atomics anywhere: bit-reproducible.
#include "fa_common.cuh"

namespace fa {

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kMaxD = 256;

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

// softmax_d = rowsum(dout * out); one warp per (token, head) row.
template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
bwd_preprocess_kernel(const Params p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y;
  const int row = blockIdx.x * kWarpsPerBlock + warp;
  const SeqInfo si = seq_info(p, b);
  if (row >= si.sq) return;
  const size_t off = ((size_t)(si.q0 + row) * p.h + h) * p.d;
  const T *o = static_cast<const T *>(p.out) + off;
  const T *g = static_cast<const T *>(p.dout) + off;
  float acc = 0.f;
  for (int c = lane * 8; c < p.d; c += 256) {
    const uint4 ro = *reinterpret_cast<const uint4 *>(o + c);
    const uint4 rg = *reinterpret_cast<const uint4 *>(g + c);
    const T *eo = reinterpret_cast<const T *>(&ro);
    const T *eg = reinterpret_cast<const T *>(&rg);
#pragma unroll
    for (int t = 0; t < 8; ++t) acc = fmaf(to_f32<T>(eo[t]), to_f32<T>(eg[t]), acc);
  }
  acc = warp_sum(acc);
  if (lane == 0) p.softmax_d[((size_t)b * p.h + h) * p.sq_rounded + row] = acc;
}

// Same result for dense inputs whose head_size / 8 is a power of two (64, 128, 256): dout and out are
// flat [token * heads][d] arrays, so consecutive lanes stream consecutive 16-byte vectors (kLanes per
// row, four rows in flight per lane) instead of one half-empty warp per row.  HBM-bound: reads dout
// and out once.
template <typename T, int kLanes>
__global__ void __launch_bounds__(256)
bwd_preprocess_flat_kernel(const T *__restrict__ o, const T *__restrict__ g, float *__restrict__ softmax_d,
                           long long rows, int h, int sq, int sq_rounded) {
  constexpr int kUnroll = 4;
  const long long nvec = rows * kLanes;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  // warp-uniform trip count (the shuffles below need the whole warp)
  for (long long v0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; v0 - lane < nvec; v0 += stride * kUnroll) {
    uint4 ro[kUnroll], rg[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long v = v0 + u * stride;
      if (v < nvec) {
        ro[u] = reinterpret_cast<const uint4 *>(o)[v];
        rg[u] = reinterpret_cast<const uint4 *>(g)[v];
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long v = v0 + u * stride;
      float acc = 0.f;
      if (v < nvec) {
        const T *eo = reinterpret_cast<const T *>(&ro[u]);
        const T *eg = reinterpret_cast<const T *>(&rg[u]);
#pragma unroll
        for (int t = 0; t < 8; ++t) acc = fmaf(to_f32<T>(eo[t]), to_f32<T>(eg[t]), acc);
      }
      // stride is a multiple of 32, so the kLanes lanes of a row stay together in one warp
#pragma unroll
      for (int off = kLanes / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
      if (v < nvec && (v % kLanes) == 0) {
        const long long r = v / kLanes;                  // (token, head) row
        const long long tok = r / h;
        const int head = (int)(r % h);
        const long long b = tok / sq;
        const int row = (int)(tok % sq);
        softmax_d[(b * h + head) * sq_rounded + row] = acc;
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
generic_bwd_dq_kernel(const Params p) {
  __shared__ float qs_all[kWarpsPerBlock][kMaxD];
  __shared__ float gs_all[kWarpsPerBlock][kMaxD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y;
  const int row = blockIdx.x * kWarpsPerBlock + warp;
  const SeqInfo si = seq_info(p, b);
  if (row >= si.sq) return;
  const int d = p.d;
  const int hk = h / (p.h / p.hk);
  const size_t roff = ((size_t)(si.q0 + row) * p.h + h) * d;
  const T *q = static_cast<const T *>(p.q) + roff;
  const T *g = static_cast<const T *>(p.dout) + roff;
  const T *kbase = static_cast<const T *>(p.k) + ((size_t)si.k0 * p.hk + hk) * d;
  const T *vbase = static_cast<const T *>(p.v) + ((size_t)si.k0 * p.hk + hk) * d;
  const size_t kv_stride = (size_t)p.hk * d;
  float *qs = qs_all[warp], *gs = gs_all[warp];
  for (int c = lane; c < d; c += 32) { qs[c] = to_f32<T>(q[c]); gs[c] = to_f32<T>(g[c]); }
  __syncwarp();

  const T *bias = nullptr;
  if (p.mask) {
    const int64_t bi = p.mask_b == 1 ? 0 : b, hi = p.mask_h == 1 ? 0 : h;
    bias = static_cast<const T *>(p.mask) + ((bi * p.mask_h + hi) * p.mask_sq + row) * p.mask_sk;
  }
  int nvis = si.sk;
  if (p.causal) nvis = min(si.sk, row + si.sk - si.sq + 1);
  if (nvis < 0) nvis = 0;

  const float lse = p.lse[((size_t)b * p.h + h) * p.sq + row];
  const float D = p.softmax_d[((size_t)b * p.h + h) * p.sq_rounded + row];
  const uint32_t bh = (uint32_t)((p.batch_base + b) * p.h + h);
  const bool drop = p.p_dropout > 0.f;
  float acc[kMaxD / 32];
#pragma unroll
  for (int t = 0; t < kMaxD / 32; ++t) acc[t] = 0.f;

  for (int kb = 0; kb < nvis; kb += 32) {
    const int key = kb + lane;
    float ds = 0.f;
    if (key < nvis && lse != INFINITY) {
      float s = dot_row<T>(qs, kbase + (size_t)key * kv_stride, d) * p.scale;
      if (bias) s += to_f32<T>(bias[key]);
      const float pn = __expf(s - lse);
      float dp = dot_row<T>(gs, vbase + (size_t)key * kv_stride, d);
      if (drop)
        dp = fa_dropout_keep(p.seed, p.offset, bh, (uint32_t)row, (uint32_t)key, p.drop_threshold)
                 ? dp * p.rp_dropout : 0.f;
      ds = pn * (dp - D) * p.scale;
    }
    const int nk = min(32, nvis - kb);
    for (int kk = 0; kk < nk; ++kk) {
      const float dsk = __shfl_sync(0xffffffffu, ds, kk);
      const T *kr = kbase + (size_t)(kb + kk) * kv_stride;
#pragma unroll
      for (int t = 0; t < kMaxD / 32; ++t) {
        const int c = lane + 32 * t;
        if (c < d) acc[t] = fmaf(dsk, to_f32<T>(kr[c]), acc[t]);
      }
    }
  }
  T *dq = static_cast<T *>(p.dq) + roff;
#pragma unroll
  for (int t = 0; t < kMaxD / 32; ++t) {
    const int c = lane + 32 * t;
    if (c < d) dq[c] = from_f32<T>(acc[t]);
  }
}

template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
generic_bwd_dkdv_kernel(const Params p) {
  __shared__ float ks_all[kWarpsPerBlock][kMaxD];
  __shared__ float vs_all[kWarpsPerBlock][kMaxD];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, hk = blockIdx.y;
  const int key = blockIdx.x * kWarpsPerBlock + warp;
  const SeqInfo si = seq_info(p, b);
  if (key >= si.sk) return;
  const int d = p.d;
  const int group = p.h / p.hk;
  const size_t koff = ((size_t)(si.k0 + key) * p.hk + hk) * d;
  const T *kr = static_cast<const T *>(p.k) + koff;
  const T *vr = static_cast<const T *>(p.v) + koff;
  float *ks = ks_all[warp], *vs = vs_all[warp];
  for (int c = lane; c < d; c += 32) { ks[c] = to_f32<T>(kr[c]); vs[c] = to_f32<T>(vr[c]); }
  __syncwarp();

  // visible query rows for this key: i >= key - (sk - sq) when causal
  int first = 0;
  if (p.causal) first = max(0, key - (si.sk - si.sq));
  const bool drop = p.p_dropout > 0.f;
  const size_t q_stride = (size_t)p.h * d;
  float dk[kMaxD / 32], dv[kMaxD / 32];
#pragma unroll
  for (int t = 0; t < kMaxD / 32; ++t) { dk[t] = 0.f; dv[t] = 0.f; }

  for (int h = hk * group; h < (hk + 1) * group; ++h) {
    const T *qbase = static_cast<const T *>(p.q) + ((size_t)si.q0 * p.h + h) * d;
    const T *gbase = static_cast<const T *>(p.dout) + ((size_t)si.q0 * p.h + h) * d;
    const float *lse = p.lse + ((size_t)b * p.h + h) * p.sq;
    const float *Dv = p.softmax_d + ((size_t)b * p.h + h) * p.sq_rounded;
    const uint32_t bh = (uint32_t)((p.batch_base + b) * p.h + h);
    const T *bias = nullptr;
    if (p.mask) {
      const int64_t bi = p.mask_b == 1 ? 0 : b, hi = p.mask_h == 1 ? 0 : h;
      bias = static_cast<const T *>(p.mask) + (bi * p.mask_h + hi) * p.mask_sq * p.mask_sk + key;
    }
    for (int rb = first; rb < si.sq; rb += 32) {
      const int row = rb + lane;
      float ds = 0.f, pd = 0.f;
      if (row < si.sq) {
        const float L = lse[row];
        if (L != INFINITY) {
          float s = dot_row<T>(ks, qbase + (size_t)row * q_stride, d) * p.scale;
          if (bias) s += to_f32<T>(bias[(size_t)row * p.mask_sk]);
          const float pn = __expf(s - L);
          float dp = dot_row<T>(vs, gbase + (size_t)row * q_stride, d);
          pd = pn;
          if (drop) {
            const bool keep = fa_dropout_keep(p.seed, p.offset, bh, (uint32_t)row, (uint32_t)key,
                                              p.drop_threshold);
            dp = keep ? dp * p.rp_dropout : 0.f;
            pd = keep ? pn * p.rp_dropout : 0.f;
          }
          ds = pn * (dp - Dv[row]) * p.scale;
        }
      }
      const int nr = min(32, si.sq - rb);
      for (int rr = 0; rr < nr; ++rr) {
        const float dsr = __shfl_sync(0xffffffffu, ds, rr);
        const float pdr = __shfl_sync(0xffffffffu, pd, rr);
        const T *qr = qbase + (size_t)(rb + rr) * q_stride;
        const T *gr = gbase + (size_t)(rb + rr) * q_stride;
#pragma unroll
        for (int t = 0; t < kMaxD / 32; ++t) {
          const int c = lane + 32 * t;
          if (c < d) {
            dk[t] = fmaf(dsr, to_f32<T>(qr[c]), dk[t]);
            dv[t] = fmaf(pdr, to_f32<T>(gr[c]), dv[t]);
          }
        }
      }
    }
  }
  T *dko = static_cast<T *>(p.dk) + koff;
  T *dvo = static_cast<T *>(p.dv) + koff;
#pragma unroll
  for (int t = 0; t < kMaxD / 32; ++t) {
    const int c = lane + 32 * t;
    if (c < d) { dko[c] = from_f32<T>(dk[t]); dvo[c] = from_f32<T>(dv[t]); }
  }
}

template <typename T>
const char *launch_bwd_t(const Params &p, cudaStream_t stream) {
  if (p.b == 0) return nullptr;
  if (p.sq > 0) {
    dim3 grid((p.sq + kWarpsPerBlock - 1) / kWarpsPerBlock, p.h, p.b);
    generic_bwd_dq_kernel<T><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p);
  }
  if (p.sk > 0) {
    dim3 grid((p.sk + kWarpsPerBlock - 1) / kWarpsPerBlock, p.hk, p.b);
    generic_bwd_dkdv_kernel<T><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p);
  }
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace

const char *launch_bwd_preprocess(const Params &p, cudaStream_t stream) {
  if (p.d % 8) return "backward: head_size must be a multiple of 8";
  if (p.b == 0 || p.sq == 0) return nullptr;
  const int lanes = p.d / 8;
  if (!p.cu_q && (lanes == 8 || lanes == 16 || lanes == 32)) {
    const long long rows = (long long)p.b * p.sq * p.h;
    const long long blocks = (rows * lanes + 256 * 4 - 1) / (256 * 4);
    const int grid = (int)(blocks < 148 * 8 ? blocks : 148 * 8);      // grid-stride, 8 CTAs per SM
#define FA_PRE(T, L)                                                                                     \
  bwd_preprocess_flat_kernel<T, L><<<grid, 256, 0, stream>>>(static_cast<const T *>(p.out),              \
                                                             static_cast<const T *>(p.dout), p.softmax_d, \
                                                             rows, p.h, p.sq, p.sq_rounded)
    if (p.is_bf16) {
      if (lanes == 8) FA_PRE(__nv_bfloat16, 8); else if (lanes == 16) FA_PRE(__nv_bfloat16, 16); else FA_PRE(__nv_bfloat16, 32);
    } else {
      if (lanes == 8) FA_PRE(__half, 8); else if (lanes == 16) FA_PRE(__half, 16); else FA_PRE(__half, 32);
    }
#undef FA_PRE
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
  }
  dim3 grid((p.sq + kWarpsPerBlock - 1) / kWarpsPerBlock, p.h, p.b);
  if (p.is_bf16)
    bwd_preprocess_kernel<__nv_bfloat16><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p);
  else
    bwd_preprocess_kernel<__half><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

const char *launch_generic_bwd(const Params &p, cudaStream_t stream) {
  if (p.d > kMaxD || p.d % 8) return "generic backward: head_size must be a multiple of 8 and <= 256";
#ifdef FA_SCALAR_BWD
  return p.is_bf16 ? launch_bwd_t<__nv_bfloat16>(p, stream) : launch_bwd_t<__half>(p, stream);
#else
  return launch_mma_bwd(p, stream);
#endif
}

}  // namespace fa
