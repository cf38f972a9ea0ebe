// fa_generic_bwd.cu -- backward preprocess, shared by both kernel families behind flash_attn_bwd /
// flash_attn_varlen_bwd (reference flash_attn.h:68-134):
//   softmax_d[b, h, i] = sum_c dout * out      (HBM-bound: reads dout and out once)
// and the entry of the all-shapes backward (fa_mma_bwd.cu).
#include "fa_common.cuh"

namespace fa {

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kMaxD = 256;

// The tcgen05 backward adds its dQ tiles into the caller's fp32 scratch, which must start at zero.  Both
// preprocess kernels zero it on the side (grid-stride 16-byte stores): one launch and no memset call.
__device__ __forceinline__ void zero_scratch(float4 *acc4, long long n4, long long tid, long long nthreads) {
  for (long long i = tid; i < n4; i += nthreads) acc4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// softmax_d = rowsum(dout * out); one warp per (token, head) row.
template <typename T>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
bwd_preprocess_kernel(const Params p, float4 *acc4, long long n4) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.z, h = blockIdx.y;
  if (acc4) {
    const long long cta = ((long long)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    zero_scratch(acc4, n4, cta * blockDim.x + threadIdx.x, (long long)gridDim.x * gridDim.y * gridDim.z * blockDim.x);
  }
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
                           long long rows, int h, int sq, int sq_rounded, float4 *acc4, long long n4) {
  constexpr int kUnroll = 4;
  if (acc4) zero_scratch(acc4, n4, blockIdx.x * (long long)blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
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

// The packed counterpart (d = 64 or 128, the fused tcgen05 backward follows): dout / out are flat [token * heads][d] arrays
// as well, only the sequence of a token has to be looked up (binary search in cu_seqlens_q: b + 1 words, cached), and
// dq_accum is zeroed ROW BY ROW next to it -- the kLanes lanes of a row clear that row's 128-byte slices of the
// [b][h][query tile][d / 32][128][32] scratch -- so that only rows that exist are written: the scratch is sized for
// max_seqlen_q, and a flat clear of it cost more than the backward's main kernel on batches of short sequences
// (128 sequences of 128 / 256 tokens: 263 of 600 us; profiles/r02m_short_end_launches.csv).  Rows past a sequence's end
// inside its last tile keep whatever they held: the main kernel adds exact zeros to them and nothing reads them.
template <typename T, int kLanes>
__global__ void __launch_bounds__(256)
bwd_preprocess_packed_kernel(const T *__restrict__ o, const T *__restrict__ g, float *__restrict__ softmax_d,
                             const int32_t *__restrict__ cu_q, int nb, int h, int sq_rounded,
                             float *__restrict__ acc, int nqt) {
  constexpr int kUnroll = 4;
  constexpr int kBlk = kLanes / 4;                       // 32-column blocks of the scratch per row (d / 32)
  const long long rows = (long long)__ldg(cu_q + nb) * h;   // total tokens: known on the device only
  const long long nvec = rows * kLanes;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
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
      float accv = 0.f;
      if (v < nvec) {
        const T *eo = reinterpret_cast<const T *>(&ro[u]);
        const T *eg = reinterpret_cast<const T *>(&rg[u]);
#pragma unroll
        for (int t = 0; t < 8; ++t) accv = fmaf(to_f32<T>(eo[t]), to_f32<T>(eg[t]), accv);
      }
#pragma unroll
      for (int off = kLanes / 2; off > 0; off >>= 1) accv += __shfl_xor_sync(0xffffffffu, accv, off);
      if (v < nvec) {
        const long long r = v / kLanes;                  // (token, head) row
        const int part = (int)(v % kLanes);
        const int tok = (int)(r / h), head = (int)(r % h);
        int lo = 0, hi = nb;                             // the sequence b with cu_q[b] <= tok < cu_q[b + 1]
        while (hi - lo > 1) {
          const int mid = (lo + hi) >> 1;
          if (__ldg(cu_q + mid) <= tok) lo = mid; else hi = mid;
        }
        const int row = tok - __ldg(cu_q + lo);
        if (part == 0) softmax_d[((long long)lo * h + head) * sq_rounded + row] = accv;
        if (acc) {                                       // my 32 bytes of the row's kBlk slices of 128 bytes
          float *p = acc + ((((long long)lo * h + head) * nqt + (row >> 7)) * kBlk + (part >> 2)) * 4096 + (row & 127) * 32 + (part & 3) * 8;
          reinterpret_cast<float4 *>(p)[0] = make_float4(0.f, 0.f, 0.f, 0.f);
          reinterpret_cast<float4 *>(p)[1] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
    }
  }
}

}  // namespace

const char *launch_bwd_preprocess(const Params &p, cudaStream_t stream, float *zero_acc, long long zero_floats) {
  if (p.d % 8) return "backward: head_size must be a multiple of 8";
  if (p.b == 0 || p.sq == 0) return nullptr;
  float4 *acc4 = reinterpret_cast<float4 *>(zero_acc);
  const long long n4 = zero_acc ? zero_floats / 4 : 0;      // the scratch is a whole number of 4096-float blocks
  const int lanes = p.d / 8;
  if (!p.cu_q && (lanes == 8 || lanes == 16 || lanes == 32)) {
    const long long rows = (long long)p.b * p.sq * p.h;
    const long long blocks = (rows * lanes + 256 * 4 - 1) / (256 * 4);
    const int grid = (int)(blocks < 148 * 8 ? blocks : 148 * 8);      // grid-stride, 8 CTAs per SM
#define FA_PRE(T, L)                                                                                     \
  bwd_preprocess_flat_kernel<T, L><<<grid, 256, 0, stream>>>(static_cast<const T *>(p.out),              \
                                                             static_cast<const T *>(p.dout), p.softmax_d, \
                                                             rows, p.h, p.sq, p.sq_rounded, acc4, n4)
    if (p.is_bf16) {
      if (lanes == 8) FA_PRE(__nv_bfloat16, 8); else if (lanes == 16) FA_PRE(__nv_bfloat16, 16); else FA_PRE(__nv_bfloat16, 32);
    } else {
      if (lanes == 8) FA_PRE(__half, 8); else if (lanes == 16) FA_PRE(__half, 16); else FA_PRE(__half, 32);
    }
#undef FA_PRE
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
  }
  if (p.cu_q && (lanes == 8 || lanes == 16)) {
    const int grid = 148 * 8;                            // grid-stride over cu_seqlens_q[b] * heads rows (a device-side number)
    const int nqt = (p.sq + 127) / 128;
#define FA_PRE(T, L)                                                                                       \
  bwd_preprocess_packed_kernel<T, L><<<grid, 256, 0, stream>>>(static_cast<const T *>(p.out),              \
                                                               static_cast<const T *>(p.dout), p.softmax_d, \
                                                               p.cu_q, p.b, p.h, p.sq_rounded, zero_acc, nqt)
    if (p.is_bf16) { if (lanes == 8) FA_PRE(__nv_bfloat16, 8); else FA_PRE(__nv_bfloat16, 16); }
    else { if (lanes == 8) FA_PRE(__half, 8); else FA_PRE(__half, 16); }
#undef FA_PRE
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
  }
  dim3 grid((p.sq + kWarpsPerBlock - 1) / kWarpsPerBlock, p.h, p.b);
  if (p.is_bf16)
    bwd_preprocess_kernel<__nv_bfloat16><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p, acc4, n4);
  else
    bwd_preprocess_kernel<__half><<<grid, kWarpsPerBlock * 32, 0, stream>>>(p, acc4, n4);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

const char *launch_generic_bwd(const Params &p, cudaStream_t stream) {
  if (p.d > kMaxD || p.d % 8) return "generic backward: head_size must be a multiple of 8 and <= 256";
  return launch_mma_bwd(p, stream);
}

}  // namespace fa
