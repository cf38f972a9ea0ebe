// fa_softmax_ret.cu -- the optional S_dmask output of the forward (reference flash_attn.h:15 `softmax_ptr`,
// :30 `return_softmax`; flash_attn_varlen_fwd :45, :58) for calls whose out / LSE come from the tcgen05 kernel.
// A second pass AFTER that kernel: P = exp(scale * Q K^T [+ bias] - LSE) from the LSE the forward has just written,
// entries the causal mask hides written as 0, entries dropout removed NEGATED -- the same
// values the all-shapes kernel returns (fa_generic_fwd.cu), so `return_softmax` no longer moves a call off the
// tcgen05 forward and no longer changes its out / LSE by a single bit.
// Debug-grade by the ABI's own intent (the output is b * h * sq_rounded * sk_rounded elements: 17 GB at config 3),
// so this is a plain warp-level MMA kernel (mma.sync m16n8k16, fp32 accumulate), not a tcgen05 pipeline: a CTA is
// (64 query rows, head, batch), its four warps hold their 16 rows of Q as A fragments for the whole walk over the
// keys, K arrives 64 keys at a time in padded shared memory.  Only [0, seqlen_q) x [0, seqlen_k) of each
// [sq_rounded, sk_rounded] plane is written, as in the all-shapes kernel.
#include "fa_common.cuh"

namespace fa {

namespace {

constexpr int kRows = 64;          // query rows per CTA (4 warps x 16)
constexpr int kKeys = 64;          // keys per step
constexpr int kThreads = 128;
constexpr float kLog2e = 1.4426950408889634f;

template <typename T> __device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <> __device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  const __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t *>(&v);
}
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t *>(&v);
}

template <typename T> struct MmaOp;
template <> struct MmaOp<__half> {
  static __device__ __forceinline__ void run(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
};
template <> struct MmaOp<__nv_bfloat16> {
  static __device__ __forceinline__ void run(float c[4], const uint32_t a[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
};

// kKS = 16-column steps of the head dimension the kernel is built for (4 / 8 / 16: head sizes up to 64 / 128 / 256;
// columns past head_size are zeros on both operands).
template <typename T, int kKS, bool kDrop, bool kBias>
__global__ void __launch_bounds__(kThreads)
softmax_ret_kernel(const Params p) {
  constexpr int kLd = kKS * 16 + 8;                       // padded row of the K tile: conflict-free fragment reads
  constexpr int kLdOut = kKeys + 8;                       // padded row of a warp's staged output block
  __shared__ __align__(16) T ks[kKeys * kLd];
  __shared__ __align__(16) T stage_all[kThreads / 32][16 * kLdOut];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;                 // fragment row / column pair of this lane
  const int b = blockIdx.z, h = blockIdx.y, row0 = blockIdx.x * kRows;
  const SeqInfo si = seq_info(p, b);
  if (row0 >= si.sq) return;
  const int d = p.d;
  const int hk = h / (p.h / p.hk);
  const T *kbase = static_cast<const T *>(p.k) + ((size_t)si.k0 * p.hk + hk) * d;
  const size_t kv_stride = (size_t)p.hk * d;
  const int shift = si.sk - si.sq;                        // causal: key n visible to row r iff n <= r + shift

  // my two rows (r[0] = fragment row g, r[1] = g + 8) and their Q fragments
  int r[2];
  float lse2[2];                                          // LSE in log2 units
  const T *bias[2] = {nullptr, nullptr};
  uint32_t qa[kKS][4];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    r[i] = row0 + warp * 16 + g + 8 * i;
    const bool live = r[i] < si.sq;
    const int rr = live ? r[i] : 0;
    lse2[i] = (live ? p.lse[((size_t)b * p.h + h) * p.sq + rr] : INFINITY) * kLog2e;
    if constexpr (kBias) {
      const int64_t bi = p.mask_b == 1 ? 0 : b, hi = p.mask_h == 1 ? 0 : h;
      bias[i] = static_cast<const T *>(p.mask) + ((bi * p.mask_h + hi) * p.mask_sq + rr) * p.mask_sk;
    }
    const T *qrow = static_cast<const T *>(p.q) + ((size_t)(si.q0 + rr) * p.h + h) * d;
#pragma unroll
    for (int s = 0; s < kKS; ++s) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {              // columns s*16 + half*8 + t4*2 + {0, 1}; d is even: a pair is in or out
        const int c = s * 16 + half * 8 + t4 * 2;
        qa[s][i + 2 * half] = (live && c < d) ? *reinterpret_cast<const uint32_t *>(qrow + c) : 0u;
      }
    }
  }
  const bool vec_store = (p.sk_rounded % 8 == 0) && (reinterpret_cast<uintptr_t>(p.softmax_ret) % 16 == 0);
  T *stage = stage_all[warp];
  T *plane = static_cast<T *>(p.softmax_ret) + ((size_t)b * p.h + h) * p.sq_rounded * (size_t)p.sk_rounded;
  const uint32_t bh = (uint32_t)((p.batch_base + b) * p.h + h);
  const float scale_log2 = p.scale * kLog2e;
  const int last_row = min(row0 + kRows, si.sq) - 1;

  for (int kb = 0; kb < si.sk; kb += kKeys) {
    const bool hidden = p.causal && kb > last_row + shift;   // (CTA-uniform) nothing of this step is visible: zeros
    float acc[kKeys / 8][4];
#pragma unroll
    for (int nt = 0; nt < kKeys / 8; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
    if (!hidden) {
      __syncthreads();                                    // the previous step's fragment reads are over
      for (int v = threadIdx.x; v < kKeys * kKS * 2; v += kThreads) {   // 16-byte vectors of the [64 keys x kKS*16] tile
        const int key = v / (kKS * 2), c = (v % (kKS * 2)) * 8;
        uint4 val = make_uint4(0u, 0u, 0u, 0u);
        if (kb + key < si.sk && c < d) val = *reinterpret_cast<const uint4 *>(kbase + (size_t)(kb + key) * kv_stride + c);
        *reinterpret_cast<uint4 *>(&ks[key * kLd + c]) = val;
      }
      __syncthreads();
#pragma unroll
      for (int s = 0; s < kKS; ++s) {
#pragma unroll
        for (int nt = 0; nt < kKeys / 8; ++nt) {
          const T *kp = &ks[(nt * 8 + g) * kLd + s * 16 + t4 * 2];
          MmaOp<T>::run(acc[nt], qa[s], *reinterpret_cast<const uint32_t *>(kp), *reinterpret_cast<const uint32_t *>(kp + 8));
        }
      }
    }
    // branch-free per entry: P = exp2(s * scale_log2 [+ bias * log2e] - lse * log2e), 0 where hidden, negated where dropped
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const int vis_end = p.causal ? min(si.sk, r[i] + shift + 1) : si.sk;   // my row sees keys [0, vis_end)
#pragma unroll
      for (int np = 0; np < kKeys / 16; ++np) {           // 16 keys = one Philox call per row
        uint32_t w[2] = {0u, 0u};                         // the random bytes of my columns in n-tiles 2np, 2np + 1
        if constexpr (kDrop) {
          if (!hidden) {
            const fa_philox4_t rnd = fa_dropout_bytes16(p.seed, p.offset, bh, (uint32_t)r[i], (uint32_t)((kb >> 4) + np));
            // columns (nt & 1) * 8 + t4 * 2 + e of the 16: word (nt & 1) * 2 + (t4 >> 1), bytes (t4 & 1) * 2 + e
            w[0] = ((t4 & 2) ? rnd.v[1] : rnd.v[0]) >> ((t4 & 1) * 16);
            w[1] = ((t4 & 2) ? rnd.v[3] : rnd.v[2]) >> ((t4 & 1) * 16);
          }
        }
#pragma unroll
        for (int sub = 0; sub < 2; ++sub) {
          const int nt = np * 2 + sub;
          const int key = kb + nt * 8 + t4 * 2;           // my two columns: key, key + 1
          float val[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            float x = acc[nt][2 * i + e] * scale_log2 - lse2[i];
            if constexpr (kBias) {
              const int col = min(key + e, si.sk - 1);    // (clamped: the value of a column past the end is discarded)
              x = fmaf(to_f32<T>(bias[i][col]), kLog2e, x);
            }
            const bool vis = key + e < vis_end;
            float v = vis ? exp2f(x) : 0.f;               // lse = +inf (no visible key, or all -inf): 0
            if constexpr (kDrop) {
              if (vis && ((w[sub] >> (8 * e)) & 0xffu) >= p.drop_threshold) v = -v;
            }
            val[e] = v;
          }
          *reinterpret_cast<uint32_t *>(&stage[(g + 8 * i) * kLdOut + nt * 8 + t4 * 2]) = pack2<T>(val[0], val[1]);
        }
      }
    }
    __syncwarp();
    // my warp's [16 rows x 64 keys] block leaves as whole 128-byte rows: lane -> (row, 16-byte vector)
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const int idx = it * 32 + lane, rw = idx >> 3, c = (idx & 7) * 8;
      const int row = row0 + warp * 16 + rw, key = kb + c;
      if (row >= si.sq || key >= si.sk) continue;
      T *out = plane + (size_t)row * p.sk_rounded + key;
      const T *src = &stage[rw * kLdOut + c];
      if (vec_store && key + 8 <= si.sk) {
        *reinterpret_cast<uint4 *>(out) = *reinterpret_cast<const uint4 *>(src);
      } else {
        for (int e = 0; e < 8 && key + e < si.sk; ++e) out[e] = src[e];
      }
    }
    __syncwarp();                                         // the block is read before the next step overwrites it
  }
}

template <typename T, bool kDrop, bool kBias>
const char *launch_t(const Params &p, cudaStream_t stream) {
  const dim3 grid((p.sq + kRows - 1) / kRows, p.h, p.b);
  if (p.d <= 64) softmax_ret_kernel<T, 4, kDrop, kBias><<<grid, kThreads, 0, stream>>>(p);
  else if (p.d <= 128) softmax_ret_kernel<T, 8, kDrop, kBias><<<grid, kThreads, 0, stream>>>(p);
  else softmax_ret_kernel<T, 16, kDrop, kBias><<<grid, kThreads, 0, stream>>>(p);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

template <typename T>
const char *launch_flags(const Params &p, cudaStream_t stream) {
  const bool drop = p.p_dropout > 0.f, bias = p.mask != nullptr;
  if (drop) return bias ? launch_t<T, true, true>(p, stream) : launch_t<T, true, false>(p, stream);
  return bias ? launch_t<T, false, true>(p, stream) : launch_t<T, false, false>(p, stream);
}

}  // namespace

const char *launch_softmax_ret(const Params &p, cudaStream_t stream) {
  if (!p.softmax_ret) return nullptr;
  if (p.d <= 0 || p.d > 256 || p.d % 8) return "softmax return: head_size must be a multiple of 8 and <= 256";
  if (p.b == 0 || p.sq == 0 || p.sk == 0) return nullptr;
  if (p.h > 65535 || p.b > 65535) return "softmax return: more than 65535 heads or batches";
  return p.is_bf16 ? launch_flags<__nv_bfloat16>(p, stream) : launch_flags<__half>(p, stream);
}

}  // namespace fa
