// fa_softmax_ret.cu -- the optional S_dmask output of the forward (reference flash_attn.h:15 `softmax_ptr`,
// :30 `return_softmax`; flash_attn_varlen_fwd :45, :58) for calls whose out / LSE come from the tcgen05 kernel.
// A second pass AFTER that kernel: P = exp(scale * Q K^T [+ bias] - LSE) from the LSE the forward has just written,
// entries the causal mask hides written as 0, entries dropout removed NEGATED -- the same
// values the all-shapes kernel returns (fa_generic_fwd.cu), so `return_softmax` no longer moves a call off the
// tcgen05 forward and no longer changes its out / LSE by a single bit.
//
// The pass is a b * h * sq * sk element WRITE (17 GB at config 3): store-bound once the scores come from tcgen05.
// One CTA per (128-query tile, head, batch), ten warps:
//   warp 8      TMA producer: the Q tile once, then the K tiles of the sequence through a ring of stages
//   warp 9      MMA issuer:   S = Q K^T (tcgen05.mma, both operands K-major in 128-byte-swizzled shared memory) into one
//               of two 128-column accumulators in tensor memory, so tile j + 1 is computed under tile j's exponentials
//   warps 0-7   TMEM lane = query row; warp w reads columns [64 (w / 4), 64 (w / 4) + 64) of its lane quadrant, turns
//               them into probabilities (one LSE per lane; the Philox draw of csrc/philox_spec.h is row-wise, one call
//               per 16 columns), parks the 16-bit tile in shared memory and the eight warps write it out as whole
//               256-byte rows.
// Key tiles the causal mask hides entirely are written as zeros without being computed.  Packed inputs: the CTA bounds its
// copy of the two tensor maps to its own sequence (ptx::patch_rows) -- rows past the end read as zero, never as memory
// behind the tensors; what a tile computes from a neighbouring sequence's rows is not written.  Only
// [0, seqlen_q) x [0, seqlen_k) of each [sq_rounded, sk_rounded] plane is written, as in the all-shapes kernel.
#include <type_traits>

#include "fa_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"

namespace fa {

namespace {

constexpr int kT = 128;            // query rows per CTA, keys per step
constexpr int kComputeWarps = 8, kLoadWarp = 8, kMmaWarp = 9;
constexpr int kThreads = 320;
constexpr float kLog2e = 1.4426950408889634f;

template <int D>
struct R {
  static constexpr uint32_t kSub = kT * 128;                  // [128 rows x 64 d] swizzled sub-tile
  static constexpr uint32_t kTile = (D / 64) * kSub;
  static constexpr int kStages = D == 256 ? 2 : (D == 128 ? 3 : 5);
  static constexpr uint32_t kQ = 0;
  static constexpr uint32_t kK = kQ + kTile;
  static constexpr uint32_t kOut = kK + kStages * kTile;      // [128 rows][16 chunks of 16 B], chunk ^= row & 7
  static constexpr uint32_t kDesc = kOut + kT * kT * 2;       // 2 x 128 B tensor-map staging (packed inputs)
  static constexpr uint32_t kBar = kDesc + 256;
  static constexpr uint32_t kNumBars = 1 + 2 * kStages + 4;
  static constexpr uint32_t kBytes = kBar + kNumBars * 8 + 16 + 1024;   // + TMEM slot + alignment slack
  static_assert(kBytes <= 232448, "over the 227 KB a CTA may ask for");
  static_assert(kBytes > 116 * 1024, "one CTA per SM is what makes the per-SM tensor-map slots private");
};

__device__ ptx::TensorMapSlot g_ret_slots[ptx::kMaxSm][2];

struct RetArgs {
  void *plane;                 // [b][h][sq_rounded][sk_rounded], 16-bit
  const float *lse;            // [b][h][sq]
  const int32_t *cu_q, *cu_k;
  const void *mask;
  int mask_b, mask_h, mask_sq, mask_sk;
  int sq, sk, sq_rounded, sk_rounded, h, hk, batch_base;   // sq / sk: dense values or packed maxima
  float scale_log2;
  uint32_t drop_threshold;
  uint64_t seed, offset;
  int causal, vec_store;
};

template <bool kBF16>
__device__ __forceinline__ uint32_t pack_pair(float lo, float hi) {
  if constexpr (kBF16) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
  } else {
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
  }
}

template <int D, bool kBF16, bool kDrop, bool kBias>
__global__ void __launch_bounds__(kThreads, 1)
tc_softmax_ret_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k, const RetArgs a) {
  using C = R<D>;
  using T = typename std::conditional<kBF16, __nv_bfloat16, __half>::type;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *sgen = smem_raw + (sbase - ptx::smem_u32(smem_raw));
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  // ---- my tile ----------------------------------------------------------------------------------------------
  const int i = blockIdx.x, head = blockIdx.y, bidx = blockIdx.z;
  int q0 = 0, k0 = 0, sq = a.sq, sk = a.sk, bcoord = bidx;
  if (a.cu_q) {
    q0 = a.cu_q[bidx]; k0 = a.cu_k[bidx];
    sq = a.cu_q[bidx + 1] - q0; sk = a.cu_k[bidx + 1] - k0;
    bcoord = 0;
  }
  if (i * kT >= sq || sk <= 0) return;                   // (the whole CTA: nothing was initialised yet)
  const int hk = head / (a.h / a.hk);
  const int shift = sk - sq;                            // causal: key n visible to row r iff n <= r + shift
  const int nkt = (sk + kT - 1) / kT;
  int n_vis = nkt;                                      // key tiles with a visible entry: the others are written as zeros
  if (a.causal) {
    const int last = min(i * kT + kT, sq) - 1 + shift;
    n_vis = last < 0 ? 0 : min(nkt, last / kT + 1);
  }

  // ---- barriers, tensor memory ------------------------------------------------------------------------------
  const uint32_t bar0 = sbase + C::kBar;
  const uint32_t q_full = bar0, k_full = bar0 + 8, k_empty = k_full + 8 * C::kStages, s_full = k_empty + 8 * C::kStages,
                 s_free = s_full + 16;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sgen + C::kBar + C::kNumBars * 8);
  if (threadIdx.x == 0) {
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < C::kStages; ++s) {
      ptx::mbar_init(k_full + 8 * s, 1);
      ptx::mbar_init(k_empty + 8 * s, 1);
    }
    for (int t = 0; t < 2; ++t) {
      ptx::mbar_init(s_full + 8 * t, 1);
      ptx::mbar_init(s_free + 8 * t, kComputeWarps * 32);
    }
    ptx::fence_mbar_init();
  }
  if (warp == kMmaWarp) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 256);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == kLoadWarp) {
    // ---- TMA producer --------------------------------------------------------------------------------------
    const CUtensorMap *mq = &tmap_q, *mk = &tmap_k;
    if (a.cu_q && n_vis > 0) {
      const uint32_t sm = ptx::smid();
      if (sm >= (uint32_t)ptx::kMaxSm) __trap();
      mq = ptx::patch_rows(tmap_q, &g_ret_slots[sm][0], sbase + C::kDesc, sgen + C::kDesc, (uint32_t)(q0 + sq), lane);
      mk = ptx::patch_rows(tmap_k, &g_ret_slots[sm][1], sbase + C::kDesc + 128, sgen + C::kDesc + 128, (uint32_t)(k0 + sk), lane);
    }
    if (lane == 0 && n_vis > 0) {
      const int qrow = (a.cu_q ? q0 : 0) + i * kT, krow = a.cu_q ? k0 : 0;   // packed: absolute token rows
      ptx::mbar_arrive_expect_tx(q_full, C::kTile);
      for (int c = 0; c < D / 64; ++c) ptx::tma_load_4d(sbase + C::kQ + c * C::kSub, mq, q_full, c * 64, head, qrow, bcoord);
      for (int jt = 0; jt < n_vis; ++jt) {
        const int st = jt % C::kStages;
        if (jt >= C::kStages) ptx::mbar_wait(k_empty + 8 * st, ((jt / C::kStages) - 1) & 1);
        ptx::mbar_arrive_expect_tx(k_full + 8 * st, C::kTile);
        for (int c = 0; c < D / 64; ++c)
          ptx::tma_load_4d(sbase + C::kK + st * C::kTile + c * C::kSub, mk, k_full + 8 * st, c * 64, hk, krow + jt * kT, bcoord);
      }
    }
  } else if (warp == kMmaWarp) {
    // ---- MMA issuer ----------------------------------------------------------------------------------------
    if (ptx::elect_one() && n_vis > 0) {
      constexpr uint32_t idesc = ptx::make_idesc(kBF16 ? ptx::kFmtBF16 : ptx::kFmtF16, ptx::kMajorK, ptx::kMajorK, kT, kT);
      constexpr uint64_t hi_k = ptx::make_smem_desc_hi(16, 1024);
      ptx::mbar_wait(q_full, 0);
      for (int jt = 0; jt < n_vis; ++jt) {
        const int st = jt % C::kStages, buf = jt & 1;
        ptx::mbar_wait(k_full + 8 * st, (jt / C::kStages) & 1);
        if (jt >= 2) ptx::mbar_wait(s_free + 8 * buf, ((jt >> 1) - 1) & 1);   // the compute warps have read S(jt - 2)
        ptx::tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          const uint32_t off = (ks / 4) * C::kSub + (ks % 4) * 32;
          ptx::mma_ss(tmem + buf * kT, ptx::smem_desc(hi_k, sbase + C::kQ + off),
                      ptx::smem_desc(hi_k, sbase + C::kK + st * C::kTile + off), idesc, ks > 0);
        }
        ptx::mma_commit(k_empty + 8 * st);
        ptx::mma_commit(s_full + 8 * buf);
      }
    }
  } else {
    // ---- compute warps ---------------------------------------------------------------------------------------
    const int qd = warp & 3, hf = warp >> 2;             // TMEM lane quadrant; which 64 of a step's 128 key columns
    const int r = qd * 32 + lane;                        // row inside the tile == TMEM lane
    const int row = i * kT + r;                          // row inside the sequence
    const bool row_ok = row < sq;
    const uint32_t lane_sel = (uint32_t)(qd * 32) << 16;
    const float lse2 = (row_ok ? a.lse[((size_t)bidx * a.h + head) * a.sq + row] : INFINITY) * kLog2e;
    const int vis_end = a.causal ? min(sk, row + shift + 1) : sk;      // my row sees keys [0, vis_end)
    const uint32_t bh = (uint32_t)((a.batch_base + bidx) * a.h + head);
    const T *bias_row = nullptr;
    if constexpr (kBias) {
      const int64_t bi = a.mask_b == 1 ? 0 : bidx, hi = a.mask_h == 1 ? 0 : head;
      bias_row = static_cast<const T *>(a.mask) + ((bi * a.mask_h + hi) * a.mask_sq + (row_ok ? row : 0)) * (int64_t)a.mask_sk;
    }
    T *plane = static_cast<T *>(a.plane) + ((size_t)bidx * a.h + head) * a.sq_rounded * (size_t)a.sk_rounded;
    uint8_t *stage = sgen + C::kOut;
    const int tid = threadIdx.x;                         // 0..255 here

    for (int jt = 0; jt < nkt; ++jt) {
      const int key0 = jt * kT + hf * 64;                // my first key column of this step
      ptx::named_bar_sync(1, kComputeWarps * 32);        // the previous step's tile has been written out of the staging
      if (jt < n_vis) {
        const int buf = jt & 1;
        ptx::mbar_wait(s_full + 8 * buf, (jt >> 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {                 // 32 columns at a time: the whole 64 would not fit beside the Philox state
          float s[32];
          uint32_t pk[16];                               // these 32 entries, two per word
          ptx::tmem_ld_x32(tmem + lane_sel + buf * kT + hf * 64 + h2 * 32, *reinterpret_cast<uint32_t(*)[32]>(&s[0]));
          ptx::tmem_wait_ld();
          if (h2 == 1) {                                 // S(jt) read: the MMA warp may overwrite it with S(jt + 2)
            ptx::tc_fence_before();
            ptx::mbar_arrive(s_free + 8 * buf);
          }
#pragma unroll
          for (int blk = 0; blk < 2; ++blk) {            // 16 columns = one Philox call of my row
            const int c0 = h2 * 32 + blk * 16;
            fa_philox4_t rnd = {{0u, 0u, 0u, 0u}};
            if constexpr (kDrop) rnd = fa_dropout_bytes16(a.seed, a.offset, bh, (uint32_t)row, (uint32_t)((key0 + c0) >> 4));
#pragma unroll
            for (int c = 0; c < 16; c += 2) {
              float v[2];
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                const int key = key0 + c0 + c + e;
                float x = fmaf(s[blk * 16 + c + e], a.scale_log2, -lse2);
                if constexpr (kBias) x = fmaf(to_f32<T>(bias_row[min(key, sk - 1)]), kLog2e, x);   // (clamped: discarded past the end)
                const bool vis = key < vis_end;
                float p = vis ? ptx::ex2(x) : 0.f;       // lse = +inf (no visible key, or all -inf): 0
                if constexpr (kDrop) {
                  const uint32_t byte = (rnd.v[(c + e) >> 2] >> (8 * ((c + e) & 3))) & 0xffu;
                  if (vis && byte >= a.drop_threshold) p = -p;
                }
                v[e] = p;
              }
              pk[(blk * 16 + c) >> 1] = pack_pair<kBF16>(v[0], v[1]);
            }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k)
            *reinterpret_cast<uint4 *>(stage + r * 256 + (((hf * 8 + h2 * 4 + k) ^ (r & 7)) << 4)) =
                make_uint4(pk[4 * k], pk[4 * k + 1], pk[4 * k + 2], pk[4 * k + 3]);
        }
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          *reinterpret_cast<uint4 *>(stage + r * 256 + (((hf * 8 + k) ^ (r & 7)) << 4)) = make_uint4(0u, 0u, 0u, 0u);
      }
      ptx::named_bar_sync(1, kComputeWarps * 32);
      // the eight warps write the tile out as whole rows: 16 lanes per 256-byte row
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int idx = it * 256 + tid, rw = idx >> 4, ch = idx & 15;
        const int grow = i * kT + rw, key = jt * kT + ch * 8;
        if (grow >= sq || key >= sk) continue;
        const uint4 val = *reinterpret_cast<const uint4 *>(stage + rw * 256 + ((ch ^ (rw & 7)) << 4));
        T *dst = plane + (size_t)grow * a.sk_rounded + key;
        if (a.vec_store && key + 8 <= sk) {
          *reinterpret_cast<uint4 *>(dst) = val;
        } else {
          const T *e = reinterpret_cast<const T *>(&val);
          for (int x = 0; x < 8 && key + x < sk; ++x) dst[x] = e[x];
        }
      }
    }
  }

  // ---- teardown -------------------------------------------------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 256);
  }
}

template <int D, bool kBF16, bool kDrop, bool kBias>
const char *launch_t(const Params &p, cudaStream_t stream) {
  using C = R<D>;
  const bool packed = p.cu_q != nullptr;
  const uint64_t nb = packed ? 1 : (uint64_t)p.b;
  CUtensorMap tq, tk;
  const char *e;
  // dense: (d, heads, seqlen, batch); packed: (d, heads, rows, 1) with the row count a placeholder each CTA replaces
  if ((e = make_tmap_rows(&tq, p.q, kBF16, p.d, p.h, (uint64_t)p.sq, nb, kT))) return e;
  if ((e = make_tmap_rows(&tk, p.k, kBF16, p.d, p.hk, (uint64_t)p.sk, nb, kT))) return e;
  RetArgs a;
  a.plane = p.softmax_ret; a.lse = p.lse; a.cu_q = p.cu_q; a.cu_k = p.cu_k;
  a.mask = p.mask; a.mask_b = (int)p.mask_b; a.mask_h = (int)p.mask_h; a.mask_sq = (int)p.mask_sq; a.mask_sk = (int)p.mask_sk;
  a.sq = p.sq; a.sk = p.sk; a.sq_rounded = p.sq_rounded; a.sk_rounded = p.sk_rounded; a.h = p.h; a.hk = p.hk;
  a.batch_base = p.batch_base;
  a.scale_log2 = p.scale * kLog2e;
  a.drop_threshold = p.drop_threshold;
  a.seed = p.seed; a.offset = p.offset;
  a.causal = p.causal ? 1 : 0;
  a.vec_store = (p.sk_rounded % 8 == 0) && (reinterpret_cast<uintptr_t>(p.softmax_ret) % 16 == 0);
  auto kern = tc_softmax_ret_kernel<D, kBF16, kDrop, kBias>;
  int dev = 0, sm_count = 0;
  if ((e = current_device(&dev, &sm_count))) return e;
  static std::atomic<uint64_t> smem_opt_in{0};   // per instantiation, one bit per device
  if ((e = ensure_dynamic_smem(kern, (int)C::kBytes, dev, smem_opt_in))) return e;
  const dim3 grid((p.sq + kT - 1) / kT, p.h, p.b);
  kern<<<grid, kThreads, C::kBytes, stream>>>(tq, tk, a);
  const cudaError_t err = cudaGetLastError();
  return err == cudaSuccess ? nullptr : cudaGetErrorString(err);
}

template <int D, bool kBF16>
const char *launch_flags(const Params &p, cudaStream_t stream) {
  const bool drop = p.p_dropout > 0.f, bias = p.mask != nullptr;
  if (drop) return bias ? launch_t<D, kBF16, true, true>(p, stream) : launch_t<D, kBF16, true, false>(p, stream);
  return bias ? launch_t<D, kBF16, false, true>(p, stream) : launch_t<D, kBF16, false, false>(p, stream);
}

}  // namespace

const char *launch_softmax_ret(const Params &p, cudaStream_t stream) {
  if (!p.softmax_ret) return nullptr;
  if (p.d <= 0 || p.d > 256 || p.d % 8) return "softmax return: head_size must be a multiple of 8 and <= 256";
  if (p.b == 0 || p.sq == 0 || p.sk == 0) return nullptr;
  if (p.h > 65535 || p.b > 65535) return "softmax return: more than 65535 heads or batches";
  if (p.d > 128) return p.is_bf16 ? launch_flags<256, true>(p, stream) : launch_flags<256, false>(p, stream);
  if (p.d > 64) return p.is_bf16 ? launch_flags<128, true>(p, stream) : launch_flags<128, false>(p, stream);
  return p.is_bf16 ? launch_flags<64, true>(p, stream) : launch_flags<64, false>(p, stream);
}

}  // namespace fa
