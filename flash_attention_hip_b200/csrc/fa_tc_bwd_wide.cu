// fa_tc_bwd_wide.cu -- FlashAttention-2 backward for sm_100a (tcgen05 + TMEM + TMA), head sizes 136..256.
// Implements flash_attn_bwd / flash_attn_varlen_bwd (reference flash_attn.h:68-99, :101-134) where
// fa_tc_bwd.cu stops: at head size 256 the dK and dV accumulators of one 128-key tile alone fill tensor
// memory (2 x 256 columns) and the four operand tiles of a 128 x 128 step alone exceed shared memory
// (4 x 64 KB).  So the OUTPUT width is split: a work item is (128-key tile, kv head, batch, half dh of the
// head dimension) and produces dK[:, dh], dV[:, dh] and its share of dQ[:, dh]; the two GEMMs that contract
// over the head dimension (S^T, dP^T) run over all 256 columns in both halves (14 MMA units per tile pair
// against the minimal 10), and the inner step is 64 queries, so K, V (resident, 2 x 64 KB) and one Q and
// one dO tile (2 x 32 KB) fit.  Per step, everything transposed as in fa_tc_bwd.cu (TMEM lane = key row):
//   (1) S^T   = K Q^T                 [128 keys x 64 q], contraction 256        -> TMEM [0, 64)
//   (2) dP^T  = V dO^T                same shape                                -> TMEM [64, 128)
//       compute warps: P^T = exp2(S^T c - LSE), dS^T = P^T o (dP^T - D)
//   (3) dV_dh += P^T dO[:, dh]        A TMEM, B smem MN-major                   -> TMEM [256, 384)
//   (4) dK_dh += dS^T Q[:, dh]        A smem K-major, B MN-major                -> TMEM [384, 512)
//   (5) dQ_dh^T = K[:, dh]^T dS^T     [128 d x 64 q]: A = K MN-major (M = d), B = dS^T MN-major
//                                                                               -> TMEM [128, 192)
// dQ is produced TRANSPOSED (an M = 64 accumulator would need the half-rate 64-lane layout): its 16 KB blocks
// [128 d][32 q] are reduce-added into dq_accum laid out [b][h][64-query tile][32-query block][d_rounded][32]
// and un-transposed by dq_convert_wide_kernel.  Q and dO are single-staged (no room for a second stage):
// the MMA order (2) (1) (3) (4) (5) lets dO(it+1) arrive under the compute phase and Q(it+1) under (5).
// Sixteen warps as in fa_tc_bwd.cu: 0-7 compute, 8 MMA issue, 9 TMA producer, 12-15 drain.
// kMode as in fa_tc_bwd.cu: 0 = the fused kernel above; num_splits == 1 and attn_mask run as two launches without
// atomics: 1 = dK / dV only (no GEMM (5), no dQ drain; with attn_mask the step's [64 q x 128 key] bias tile arrives by
// TMA in the then idle 16 KB staging block), 2 = dQ only, MIRRORED: an item is (128-QUERY tile, query head, batch, dh);
// Q and dO are resident (the launcher passes their tensor maps in the K / V positions), K and V stream in 64-KEY steps,
// so (1) S = Q K^T, (2) dP = dO V^T have TMEM lane = query row and (4) accumulates dQ[:, dh] = dS K[:, dh] in the dK
// columns; LSE / D are two scalars per lane, mask and dropout are row-wise.
#include "fa_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"
#include "fa_bwd_dev.cuh"

namespace fa {

namespace {

using namespace bwd_dev;

constexpr int kT = 128;            // keys per work item
constexpr int kQ = 64;             // queries per inner step
constexpr int kThreads = 512;
constexpr int kMmaWarp = 8, kLoadWarp = 9, kDrainWarp0 = 12;
constexpr float kLog2e = 1.4426950408889634f;

struct W {
  static constexpr uint32_t kSubK = kT * 128;          // [128 rows x 64 d] swizzled sub-tile of K / V
  static constexpr uint32_t kSubQ = kQ * 128;          // [64 rows x 64 d] sub-tile of Q / dO
  static constexpr uint32_t kK = 0;                    // 4 sub-tiles
  static constexpr uint32_t kV = kK + 4 * kSubK;
  static constexpr uint32_t kQs = kV + 4 * kSubK;      // one stage
  static constexpr uint32_t kDO = kQs + 4 * kSubQ;     // one stage
  static constexpr uint32_t kDS = kDO + 4 * kSubQ;     // dS^T [128 keys x 64 q] 16-bit = 16 KB
  static constexpr uint32_t kStage = kDS + 16384;      // one 16 KB staging block (dQ blocks, dK / dV boxes)
  static constexpr uint32_t kStat = kStage + 16384;    // 2 x {lse2[64], D[64]} fp32
  static constexpr uint32_t kBar = kStat + 2 * 512;
  static constexpr uint32_t kNumBars = 20;
  static constexpr uint32_t kDesc = kBar + 256;        // 6 x 128 B tensor-map staging (packed inputs)
  static constexpr uint32_t kBytes = kDesc + 6 * 128;
  static_assert(kBytes <= 232448, "over the 227 KB a CTA may ask for");
};

__device__ SchedSlot g_bwdw_sched[kSchedSlots];
__device__ ptx::TensorMapSlot g_bwdw_slots[ptx::kMaxSm][2][6];

struct WideArgs {
  float *dq_accum;
  const float *lse, *softmax_d;
  const int32_t *cu_q, *cu_k;
  int b, sq, sk, sq_rounded, h, hk, nq64, batch_base;   // sq / sk / nq64: dense values or packed maxima
  int d, d_r;            // head_size; head_size rounded up to 32 (rows per 32-query block of dq_accum)
  float scale, scale_log2, rp_dropout, keep_prob;
  uint32_t drop_threshold, key0, key1, ctr3;
  int causal;
  const void *mask;      // additive bias [mask_b, mask_h, mask_sq, mask_sk] (modes 1 and 2 only), or null
  int mask_b, mask_h, mask_sq, mask_sk;
  int mask_tma;          // mode 1: bias tiles arrive by TMA (tmap_bias)
  int hk_kv;             // K/V heads (mode 2 overwrites hk with the query head count)
  int nkt, n_items;      // key tiles per sequence; items = 2 * nkt * hk * b   (mode 2: nkt = 128-query tiles)
  int persistent;
  SchedSlot *sched;
};

template <bool kBF16, bool kDropout, bool kMask, int kMode>
__global__ void __launch_bounds__(kThreads, 1)
tc_bwd_wide_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
                   const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_do,
                   const __grid_constant__ CUtensorMap tmap_dk, const __grid_constant__ CUtensorMap tmap_dv,
                   const __grid_constant__ CUtensorMap tmap_bias, const WideArgs a) {
  // mirrored (mode 2): tmap_k / tmap_v are the maps of Q / dO (128-row boxes), tmap_q / tmap_do those of K / V (64-row
  // boxes), tmap_dk is dq's: the slots, GEMMs and barriers below keep their names
  constexpr bool kMirror = kMode == 2;
  constexpr bool kBiasTma = kMask && kMode == 1;
  static_assert(kMode != 0 || !kMask, "attn_mask runs as the two-launch pair");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t sbase = ptx::smem_u32(smem_raw);
  if (sbase & 1023u) __trap();
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  struct Item {
    int dh, j, hk, bidx, sq, sk, qrow0, krow0, bcoord, shift, i_min, n_i, n_it;
    bool valid;
  };
  const int group_real = a.h / a.hk_kv;
  const int group = kMirror ? 1 : group_real;          // mirrored: `hk` IS the query head
  auto decode = [&](int L) {
    Item I;
    I.dh = L & 1; L >>= 1;                      // the two halves of one key tile run side by side: shared L2 lines
    I.j = L % a.nkt; I.hk = (L / a.nkt) % a.hk; I.bidx = L / (a.nkt * a.hk);
    if (kMirror && a.causal) I.j = a.nkt - 1 - I.j;    // the last query tile sees the most keys: heaviest first
    I.sq = a.sq; I.sk = a.sk; I.qrow0 = 0; I.krow0 = 0; I.bcoord = I.bidx;
    if (a.cu_q) {
      I.qrow0 = a.cu_q[I.bidx]; I.krow0 = a.cu_k[I.bidx];
      I.sq = a.cu_q[I.bidx + 1] - I.qrow0; I.sk = a.cu_k[I.bidx + 1] - I.krow0;
      I.bcoord = 0;
    }
    I.shift = I.sk - I.sq;                       // causal: key n visible to row r iff n <= r + shift
    if constexpr (kMirror) {                     // `j` = my 128-QUERY tile; the loop walks the 64-key steps its rows see
      I.valid = I.j * kT < I.sq;
      int nk = (I.sk + kQ - 1) / kQ;
      if (a.causal) {
        const int last_key = I.j * kT + kT - 1 + I.shift;
        nk = last_key < 0 ? 0 : min(nk, last_key / kQ + 1);
      }
      I.i_min = 0; I.n_i = nk; I.n_it = nk;
      return I;
    }
    I.valid = I.j * kT < I.sk;
    const int nq = (I.sq + kQ - 1) / kQ;
    I.i_min = 0;
    if (a.causal) {
      const int first_row = I.j * kT - I.shift;
      I.i_min = first_row <= 0 ? 0 : first_row / kQ;
    }
    I.n_i = max(0, nq - I.i_min);
    I.n_it = I.n_i * group;
    return I;
  };
  if (!a.persistent && !decode((int)blockIdx.x).valid) return;

  const uint32_t bar0 = sbase + W::kBar;
  const uint32_t b_kv_full = bar0, b_q_full = bar0 + 8, b_q_empty = bar0 + 16, b_do_full = bar0 + 24,
                 b_do_empty = bar0 + 32, b_s_full = bar0 + 40, b_dp_full = bar0 + 48, b_p_ready = bar0 + 56,
                 b_ds_ready = bar0 + 64, b_dq_full = bar0 + 72, b_dq_drained = bar0 + 80, b_acc_done = bar0 + 88,
                 b_epi_tmem_free = bar0 + 96, b_sched_full = bar0 + 104 /*[2]*/, b_sched_empty = bar0 + 120 /*[2]*/,
                 b_bias_full = bar0 + 136, b_bias_empty = bar0 + 144, b_epi_smem_free = bar0 + 152;
  static_assert(W::kNumBars >= 20 && W::kNumBars * 8 + 32 <= 256, "barrier block too small");
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem_raw + W::kBar + W::kNumBars * 8);
  volatile int *sched_item = reinterpret_cast<volatile int *>(tmem_slot + 2);   // [2]

  if (threadIdx.x == 0) {
    ptx::mbar_init(b_kv_full, 1);
    ptx::mbar_init(b_q_full, 2);          // TMA expect_tx arrive + stats arrive
    ptx::mbar_init(b_q_empty, 1);
    ptx::mbar_init(b_do_full, 1);     ptx::mbar_init(b_do_empty, 1);
    ptx::mbar_init(b_s_full, 1);      ptx::mbar_init(b_dp_full, 1);
    ptx::mbar_init(b_p_ready, 256);   ptx::mbar_init(b_ds_ready, 256);
    ptx::mbar_init(b_dq_full, 1);     ptx::mbar_init(b_dq_drained, 128);
    ptx::mbar_init(b_acc_done, 1);    ptx::mbar_init(b_epi_tmem_free, 128);
    ptx::mbar_init(b_bias_full, 1);   ptx::mbar_init(b_bias_empty, 256);   ptx::mbar_init(b_epi_smem_free, 1);
    for (int t = 0; t < 2; ++t) {
      ptx::mbar_init(b_sched_full + 8 * t, 1);
      ptx::mbar_init(b_sched_empty + 8 * t, 1 + 128 + 256);    // MMA lane, drain warps, compute warps
    }
    ptx::fence_mbar_init();
  }
  if (warp == kLoadWarp && lane == 0) {
    ptx::prefetch_tensormap(&tmap_q);  ptx::prefetch_tensormap(&tmap_k);
    ptx::prefetch_tensormap(&tmap_v);  ptx::prefetch_tensormap(&tmap_do);
  }
  if (warp == kMmaWarp) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem, tDP = tmem + 64, tDQ = tmem + 128, tDV = tmem + 256, tDK = tmem + 384;

  auto next_item = [&](int n) {
    const int slot = n & 1;
    ptx::mbar_wait(b_sched_full + 8 * slot, (n >> 1) & 1);
    const int L = sched_item[slot];
    ptx::mbar_arrive(b_sched_empty + 8 * slot);
    return L;
  };

  if (warp >= 8 && warp < 12) {
    if (warp == kLoadWarp) {
      // =========================== TMA producer; all 32 lanes stage LSE / D ========================
      const CUtensorMap *mq = &tmap_q, *mk = &tmap_k, *mv = &tmap_v, *mdo = &tmap_do;
      int mapped_b = -1, set = 0, nv = 0, g0 = 0;
      for (int n = 0;; ++n) {
        int L = 0;
        if (lane == 0) {
          const int slot = n & 1;
          if (n >= 2) ptx::mbar_wait(b_sched_empty + 8 * slot, ((n >> 1) - 1) & 1);
          L = a.persistent ? sched_next_item(a.sched, a.n_items) : (n == 0 ? (int)blockIdx.x : -1);
          sched_item[slot] = L;
          ptx::mbar_arrive(b_sched_full + 8 * slot);
        }
        L = __shfl_sync(0xffffffffu, L, 0);
        if (L < 0) break;
        const Item I = decode(L);
        if (!I.valid) continue;
        if (nv >= 1) ptx::mbar_wait(b_acc_done, (nv - 1) & 1);   // the previous item's MMAs have read K and V
        if (a.cu_q && I.bidx != mapped_b) {
          // packed inputs: rows past this sequence's end must read as zero, not as the next sequence (fa_tc_bwd.cu)
          const uint32_t sm = ptx::smid();
          if (sm >= (uint32_t)ptx::kMaxSm) __trap();
          const uint32_t d0 = sbase + W::kDesc;
          uint8_t *gen0 = smem_raw + W::kDesc;
          // (mirrored: the maps in the q / do positions are K's and V's, those in the k / v positions Q's and dO's)
          const uint32_t str_rows = kMirror ? (uint32_t)(I.krow0 + I.sk) : (uint32_t)(I.qrow0 + I.sq);
          const uint32_t res_rows = kMirror ? (uint32_t)(I.qrow0 + I.sq) : (uint32_t)(I.krow0 + I.sk);
          mq = ptx::patch_rows(tmap_q, &g_bwdw_slots[sm][set][0], d0, gen0, str_rows, lane);
          mdo = ptx::patch_rows(tmap_do, &g_bwdw_slots[sm][set][1], d0 + 128, gen0 + 128, str_rows, lane);
          mk = ptx::patch_rows(tmap_k, &g_bwdw_slots[sm][set][2], d0 + 256, gen0 + 256, res_rows, lane);
          mv = ptx::patch_rows(tmap_v, &g_bwdw_slots[sm][set][3], d0 + 384, gen0 + 384, res_rows, lane);
          mapped_b = I.bidx;
          set ^= 1;
        }
        const int res_row0 = kMirror ? I.qrow0 : I.krow0, str_row0 = kMirror ? I.krow0 : I.qrow0;
        if (lane == 0) {
          ptx::mbar_arrive_expect_tx(b_kv_full, 8 * W::kSubK);
          for (int c = 0; c < 4; ++c) {
            ptx::tma_load_4d(sbase + W::kK + c * W::kSubK, mk, b_kv_full, c * 64, I.hk, res_row0 + I.j * kT, I.bcoord);
            ptx::tma_load_4d(sbase + W::kV + c * W::kSubK, mv, b_kv_full, c * 64, I.hk, res_row0 + I.j * kT, I.bcoord);
          }
        }
        for (int it = 0; it < I.n_it; ++it) {
          const int hq = I.hk * group + it / I.n_i, i = I.i_min + it % I.n_i;
          const int str_head = kMirror ? I.hk / group_real : hq;
          const int g = g0 + it;
          if constexpr (kBiasTma) {
            if (a.mask_tma) {
              if (g >= 1) ptx::mbar_wait(b_bias_empty, (g - 1) & 1);                  // the compute warps have read tile g - 1
              if (it == 0 && nv >= 1) ptx::mbar_wait(b_epi_smem_free, (nv - 1) & 1);  // the last item's dK / dV have left the staging
              if (lane == 0) {
                ptx::mbar_arrive_expect_tx(b_bias_full, kT * kQ * 2);
                ptx::tma_load_4d(sbase + W::kStage, &tmap_bias, b_bias_full, I.j * kT, i * kQ, a.mask_h == 1 ? 0 : hq,
                                 a.mask_b == 1 ? 0 : I.bidx);
              }
            }
          }
          // dO first: its slot frees after (3), long before Q's does
          if (g >= 1) ptx::mbar_wait(b_do_empty, (g - 1) & 1);
          if (lane == 0) {
            ptx::mbar_arrive_expect_tx(b_do_full, 4 * W::kSubQ);
            for (int c = 0; c < 4; ++c)
              ptx::tma_load_4d(sbase + W::kDO + c * W::kSubQ, mdo, b_do_full, c * 64, str_head, str_row0 + i * kQ, I.bcoord);
          }
          if (g >= 1) ptx::mbar_wait(b_q_empty, (g - 1) & 1);
          if (lane == 0) {
            ptx::mbar_arrive_expect_tx(b_q_full, 4 * W::kSubQ);
            for (int c = 0; c < 4; ++c)
              ptx::tma_load_4d(sbase + W::kQs + c * W::kSubQ, mq, b_q_full, c * 64, str_head, str_row0 + i * kQ, I.bcoord);
          }
          if constexpr (kMirror) {
            if (lane == 0) ptx::mbar_arrive(b_q_full);     // (LSE / D are per lane here: nothing to stage)
          } else {
            // staged NEGATED: -LSE * log2(e) (-inf past the sequence end => P = 0 there), -D (1 - p)
            const float *lse = a.lse + ((size_t)I.bidx * a.h + hq) * a.sq;
            const float *dd = a.softmax_d + ((size_t)I.bidx * a.h + hq) * a.sq_rounded;
            const uint32_t st = sbase + W::kStat + (g & 1) * 512;
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              const int r = t * 32 + lane, row = i * kQ + r;
              const float l = row < I.sq ? lse[row] : INFINITY;
              const float dv = row < I.sq ? dd[row] : 0.f;
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(st + 4 * r), "f"(-l * kLog2e) : "memory");
              asm volatile("st.shared.f32 [%0], %1;" ::"r"(st + 256 + 4 * r), "f"(-dv * a.keep_prob) : "memory");
            }
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(b_q_full);
          }
        }
        if constexpr (kBiasTma) {    // (every phase of epi_smem_free is waited for, in order)
          if (a.mask_tma && I.n_it == 0 && nv >= 1) ptx::mbar_wait(b_epi_smem_free, (nv - 1) & 1);
        }
        g0 += I.n_it;
        ++nv;
      }
    } else if (warp == kMmaWarp && ptx::elect_one()) {
      // =========================== MMA issuer =======================================================
      constexpr uint32_t fmt = kBF16 ? ptx::kFmtBF16 : ptx::kFmtF16;
      constexpr uint32_t id_s = ptx::make_idesc(fmt, ptx::kMajorK, ptx::kMajorK, kT, kQ);        // (1) (2)
      constexpr uint32_t id_acc = ptx::make_idesc(fmt, ptx::kMajorK, ptx::kMajorMN, kT, 128);    // (3) (4)
      constexpr uint32_t id_dq = ptx::make_idesc(fmt, ptx::kMajorMN, ptx::kMajorMN, 128, kQ);    // (5)
      constexpr uint64_t hi_k = ptx::make_smem_desc_hi(16, 1024);
      constexpr uint64_t hi_mn_q = ptx::make_smem_desc_hi(W::kSubQ, 1024);   // Q / dO / dS^T read MN-major
      constexpr uint64_t hi_mn_k = ptx::make_smem_desc_hi(W::kSubK, 1024);   // K read MN-major (M = d)
      const int nks = (a.d + 15) / 16;     // k-steps of the two GEMMs that contract over the head dimension (the rest is zero fill)
      int g0 = 0, nv = 0;
      for (int n = 0;; ++n) {
        const int L = next_item(n);
        if (L < 0) break;
        const Item I = decode(L);
        if (!I.valid) continue;
        const uint32_t half_q = 2 * I.dh * W::kSubQ, half_k = 2 * I.dh * W::kSubK;
        ptx::mbar_wait(b_kv_full, nv & 1);
        ptx::tc_fence_after();
        // An item without steps (mirrored pass: a query tile that sees no key) commits acc_done at once.  The drain warps wait
        // for acc_done by PARITY: were this warp to run two items ahead of them, they would wait for the wrong phase (seen as a
        // hang on a packed causal batch with sk << sq: runs of such items).  So an empty item first waits for the previous
        // item's epilogue, which a non-empty one does before its first accumulating MMA anyway.
        if (I.n_it == 0 && nv >= 1) ptx::mbar_wait(b_epi_tmem_free, (nv - 1) & 1);
        for (int it = 0; it < I.n_it; ++it) {
          const int g = g0 + it;
          const uint32_t b16 = ptx::opaque(sbase >> 4);
          ptx::mbar_wait(b_do_full, g & 1);
          ptx::tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < 16; ++ks)                                      // (2) dP^T = V dO^T
            if (ks < nks) ptx::mma_ss(tDP, ptx::smem_desc_at(hi_k, b16, W::kV + (ks / 4) * W::kSubK + (ks % 4) * 32),
                        ptx::smem_desc_at(hi_k, b16, W::kDO + (ks / 4) * W::kSubQ + (ks % 4) * 32), id_s, ks > 0);
          ptx::mma_commit(b_dp_full);
          if constexpr (kMirror) ptx::mma_commit(b_do_empty);                  // no (3) here: the dO slot is free after (2)
          ptx::mbar_wait(b_q_full, g & 1);
          ptx::tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < 16; ++ks)                                      // (1) S^T = K Q^T
            if (ks < nks) ptx::mma_ss(tS, ptx::smem_desc_at(hi_k, b16, W::kK + (ks / 4) * W::kSubK + (ks % 4) * 32),
                        ptx::smem_desc_at(hi_k, b16, W::kQs + (ks / 4) * W::kSubQ + (ks % 4) * 32), id_s, ks > 0);
          ptx::mma_commit(b_s_full);
          ptx::mbar_wait(b_p_ready, g & 1);
          // the first dV / dK MMAs of an item overwrite the accumulators: the previous item's epilogue must have read them
          if (it == 0 && nv >= 1) ptx::mbar_wait(b_epi_tmem_free, (nv - 1) & 1);
          ptx::tc_fence_after();
          if constexpr (!kMirror) {
#pragma unroll
            for (int ks = 0; ks < kQ / 16; ++ks)                               // (3) dV_dh += P^T dO[:, dh]
              ptx::mma_ts(tDV, tS + (ks / 2) * 32 + (ks % 2) * 8,
                          ptx::smem_desc_at(hi_mn_q, b16 + (half_q >> 4), W::kDO + ks * 2048), id_acc, it > 0 || ks > 0);
            ptx::mma_commit(b_do_empty);
          }
          ptx::mbar_wait(b_ds_ready, g & 1);
          ptx::tc_fence_after();
#pragma unroll
          for (int ks = 0; ks < kQ / 16; ++ks)                                 // (4) dK_dh += dS^T Q[:, dh]
            ptx::mma_ss(tDK, ptx::smem_desc_at(hi_k, b16, W::kDS + ks * 32),
                        ptx::smem_desc_at(hi_mn_q, b16 + (half_q >> 4), W::kQs + ks * 2048), id_acc, it > 0 || ks > 0);
          ptx::mma_commit(b_q_empty);
          if constexpr (kMode == 0) {
            if (g > 0) {
              ptx::mbar_wait(b_dq_drained, (g - 1) & 1);
              ptx::tc_fence_after();
            }
#pragma unroll
            for (int ks = 0; ks < kT / 16; ++ks)                               // (5) dQ_dh^T = K[:, dh]^T dS^T
              ptx::mma_ss(tDQ, ptx::smem_desc_at(hi_mn_k, b16 + (half_k >> 4), W::kK + ks * 2048),
                          ptx::smem_desc_at(hi_mn_q, b16, W::kDS + ks * 2048), id_dq, ks > 0);
            ptx::mma_commit(b_dq_full);
          }
        }
        ptx::mma_commit(b_acc_done);
        g0 += I.n_it;
        ++nv;
      }
    }
  } else if (warp >= kDrainWarp0) {
    // =========================== drain warps: dQ^T blocks, then dK / dV of the item =================
    const int dw = warp - kDrainWarp0;
    const int r = dw * 32 + lane;                      // TMEM lane: d row of dQ^T, key row of dK / dV
    const uint32_t lane_sel = (uint32_t)(dw * 32) << 16;
    const bool leader = (warp == kDrainWarp0 && lane == 0);
    const uint32_t buf = sbase + W::kStage;
    const CUtensorMap *mdk = &tmap_dk, *mdv = &tmap_dv;
    int g0 = 0, nv = 0, mapped_b = -1, set = 0;
    for (int n = 0;; ++n) {
      const int L = next_item(n);
      if (L < 0) break;
      const Item I = decode(L);
      if (!I.valid) continue;
      if (a.cu_q && warp == kDrainWarp0 && I.bidx != mapped_b) {
        if (lane == 0) ptx::tma_store_wait<0>();
        __syncwarp();
        const uint32_t sm = ptx::smid();
        if (sm >= (uint32_t)ptx::kMaxSm) __trap();
        const uint32_t d0 = sbase + W::kDesc;
        uint8_t *gen0 = smem_raw + W::kDesc;
        const uint32_t out_rows = kMirror ? (uint32_t)(I.qrow0 + I.sq) : (uint32_t)(I.krow0 + I.sk);   // mirrored: the map is dq's
        mdk = ptx::patch_rows(tmap_dk, &g_bwdw_slots[sm][set][4], d0 + 512, gen0 + 512, out_rows, lane);
        mdv = ptx::patch_rows(tmap_dv, &g_bwdw_slots[sm][set][5], d0 + 640, gen0 + 640, out_rows, lane);
        mapped_b = I.bidx;
        set ^= 1;
      }
      // d rows of this half that exist in dq_accum (head_size rounded up to 32)
      const int rows_valid = min(128, a.d_r - I.dh * 128);
      for (int it = 0; kMode == 0 && it < I.n_it; ++it) {
        const int hq = I.hk * group + it / I.n_i, i = I.i_min + it % I.n_i;
        // dq_accum: [b][h][64-query tile][32-query block][d_r rows][32 floats], 16-byte chunks XOR-swizzled by row
        float *gdst = a.dq_accum + ((((size_t)I.bidx * a.h + hq) * a.nq64 + i) * 2 * a.d_r + I.dh * 128) * 32;
        ptx::mbar_wait(b_dq_full, (g0 + it) & 1);
        ptx::tc_fence_after();
        uint32_t v[2][32];
        ptx::tmem_ld_x32(tDQ + lane_sel, v[0]);
        ptx::tmem_ld_x32(tDQ + lane_sel + 32, v[1]);
        ptx::tmem_wait_ld();
        ptx::tc_fence_before();
        ptx::mbar_arrive(b_dq_drained);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (leader) ptx::tma_store_wait_read<0>();   // the block that used the staging buffer has been read
          ptx::named_bar_sync(1, 128);
#pragma unroll
          for (int k = 0; k < 8; ++k)
            st_shared_v4(buf + r * 128 + ((k ^ (r & 7)) << 4), v[c][4 * k], v[c][4 * k + 1], v[c][4 * k + 2], v[c][4 * k + 3]);
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(1, 128);
          if (leader) {
            if (rows_valid > 0) bulk_reduce_add_f32(gdst + (size_t)c * a.d_r * 32, buf, (uint32_t)rows_valid * 128u);
            ptx::tma_store_commit();
          }
        }
      }
      // ---- epilogue: dK_dh, dV_dh -> global through bulk tensor stores (clipped by the maps) ----
      ptx::mbar_wait(b_acc_done, nv & 1);
      ptx::tc_fence_after();
#pragma unroll
      for (int which = kMirror ? 1 : 0; which < 2; ++which) {      // mirrored: only the dK columns are used (they hold dQ)
        const float mul = which ? a.scale * a.rp_dropout : a.rp_dropout;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (leader) ptx::tma_store_wait_read<0>();
          ptx::named_bar_sync(1, 128);
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            uint32_t v[32];
            if (I.n_it > 0) {
              ptx::tmem_ld_x32((which ? tDK : tDV) + lane_sel + c * 64 + hf * 32, v);
              ptx::tmem_wait_ld();
            } else {
#pragma unroll
              for (int x = 0; x < 32; ++x) v[x] = 0;
            }
            if (which == 1 && c == 1 && hf == 1) {       // dK / dV read: the next item's MMAs may overwrite them
              ptx::tc_fence_before();
              ptx::mbar_arrive(b_epi_tmem_free);
            }
#pragma unroll
            for (int x = 0; x < 4; ++x)
              st_shared_v4(buf + r * 128 + (((hf * 4 + x) ^ (r & 7)) << 4),
                           pack2(__uint_as_float(v[8 * x]) * mul, __uint_as_float(v[8 * x + 1]) * mul, kBF16),
                           pack2(__uint_as_float(v[8 * x + 2]) * mul, __uint_as_float(v[8 * x + 3]) * mul, kBF16),
                           pack2(__uint_as_float(v[8 * x + 4]) * mul, __uint_as_float(v[8 * x + 5]) * mul, kBF16),
                           pack2(__uint_as_float(v[8 * x + 6]) * mul, __uint_as_float(v[8 * x + 7]) * mul, kBF16));
          }
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(1, 128);
          if (leader) {
            const int dcol = I.dh * 128 + c * 64;
            if (dcol < a.d) ptx::tma_store_4d(which ? mdk : mdv, buf, dcol, I.hk, (kMirror ? I.qrow0 : I.krow0) + I.j * kT, I.bcoord);
            ptx::tma_store_commit();
          }
        }
      }
      if constexpr (kBiasTma) {
        if (a.mask_tma && leader) {        // the staging block has been read by the stores: the next item's bias tile may land
          ptx::tma_store_wait_read<0>();
          ptx::mbar_arrive(b_epi_smem_free);
        }
      }
      g0 += I.n_it;
      ++nv;
    }
    if (leader) ptx::tma_store_wait<0>();
  } else {
    // =========================== compute warps ======================================================
    const int w = warp >> 2;                           // which 32 query columns
    const int nrow = (warp & 3) * 32 + lane;           // key row inside the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    const float c2 = a.scale_log2;
    uint32_t add7 = 0;
    bool t_msb = false;
    if constexpr (kDropout) {
      add7 = (0x80u - (a.drop_threshold & 0x7fu)) * 0x01010101u;
      t_msb = a.drop_threshold >= 128u;
    }
    int g0 = 0;
    for (int item_n = 0;; ++item_n) {
      const int L = next_item(item_n);
      if (L < 0) break;
      const Item I = decode(L);
      if (!I.valid) continue;
      const int n_abs = I.j * kT + nrow;                 // my key -- mirrored: my QUERY row
      const bool kv_valid = n_abs < (kMirror ? I.sq : I.sk);
      const bool ragged_k = I.j * kT + kT > I.sk;
      float nl_mine = -INFINITY, nd_mine = 0.f;          // mirrored: LSE / D of my own row, negated and scaled like the staged ones
      if constexpr (kMirror) {
        if (kv_valid) {
          nl_mine = -a.lse[((size_t)I.bidx * a.h + I.hk) * a.sq + n_abs] * kLog2e;
          nd_mine = -a.softmax_d[((size_t)I.bidx * a.h + I.hk) * a.sq_rounded + n_abs] * a.keep_prob;
        }
      }
      for (int it = 0; it < I.n_it; ++it) {
        const int hq = I.hk * group + it / I.n_i, i = I.i_min + it % I.n_i;
        const int g = g0 + it;
        const uint32_t stat = sbase + W::kStat + (g & 1) * 512 + w * 128;   // lse2 / D of my 32 query columns
        const int q_base = i * kQ + w * 32;

        // dropout keep bits of my 32 (query, key) pairs (csrc/philox_spec.h): the lanes of a half-warp share the
        // 16-key block, so each lane draws the 16 bytes of ONE query row per 16-row block (two calls) and the
        // half-warp exchanges them with 16x16 bit transposes.  Drawn before the wait for S^T (they do not depend on it).
        uint32_t keep = 0xffffffffu;                   // bit t: keep(query q_base + t, my key)
        if constexpr (kDropout) {
          const uint32_t bh = (uint32_t)((a.batch_base + I.bidx) * a.h + hq);
          uint32_t w01 = 0;
#pragma unroll
          for (int blk = 0; blk < 2; ++blk) {
            // mirrored: my row, the two 16-key blocks of my 32 columns (the forward's own mapping, no exchange)
            const fa_philox4_t rnd = kMirror
                ? fa_philox4x32((uint32_t)n_abs, (uint32_t)((q_base + blk * 16) >> 4), bh, a.ctr3, a.key0, a.key1)
                : fa_philox4x32((uint32_t)(q_base + blk * 16 + (lane & 15)), (uint32_t)(n_abs >> 4), bh, a.ctr3, a.key0, a.key1);
            w01 |= keep_bits16(rnd, add7, t_msb) << (16 * blk);
          }
          if constexpr (!kMirror) {
            bit_transpose_stage1<8>(w01, lane);
            bit_transpose_stage1<4>(w01, lane);
            bit_transpose_stage1<2>(w01, lane);
            bit_transpose_stage1<1>(w01, lane);
          }
          keep = w01;
        }

        // attn_mask: my 32 bias values, fetched before the wait for S (they do not depend on it)
        uint16_t mbias[kMask ? 32 : 1];
        if constexpr (kMask && kMirror) {        // my row's 32 keys: 64 contiguous bytes
          const uint16_t *mp = static_cast<const uint16_t *>(a.mask) +
              (((size_t)(a.mask_b == 1 ? 0 : I.bidx) * a.mask_h + (a.mask_h == 1 ? 0 : hq)) * a.mask_sq + n_abs) * (size_t)a.mask_sk + q_base;
          const bool fast = kv_valid && q_base + 32 <= I.sk && a.mask_sk % 8 == 0 && (reinterpret_cast<uintptr_t>(a.mask) & 15) == 0;
          if (fast) {
#pragma unroll
            for (int v8 = 0; v8 < 4; ++v8) {
              const uint4 x = __ldg(reinterpret_cast<const uint4 *>(mp) + v8);
              const uint32_t w4[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
              for (int e = 0; e < 8; ++e) mbias[8 * v8 + e] = (uint16_t)(w4[e / 2] >> (16 * (e & 1)));
            }
          } else {
#pragma unroll
            for (int c = 0; c < 32; ++c) mbias[c] = (kv_valid && q_base + c < I.sk) ? __ldg(mp + c) : (uint16_t)0;
          }
        } else if (kBiasTma && a.mask_tma) {     // the step's [64 query][128 key] tile in the staging block
          ptx::mbar_wait(b_bias_full, g & 1);
          const uint16_t *bt = reinterpret_cast<const uint16_t *>(smem_raw + W::kStage) + (size_t)(w * 32) * kT + nrow;
          const bool rows_ragged = q_base + 32 > I.sq;
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const uint16_t vb = bt[c * kT];
            mbias[c] = (kv_valid && (!rows_ragged || q_base + c < I.sq)) ? vb : (uint16_t)0;   // (padding may hold anything)
          }
          ptx::mbar_arrive(b_bias_empty);
        } else if constexpr (kMask) {            // my key COLUMN of the bias: one two-byte load per query row
          const uint16_t *mp = static_cast<const uint16_t *>(a.mask) +
              (((size_t)(a.mask_b == 1 ? 0 : I.bidx) * a.mask_h + (a.mask_h == 1 ? 0 : hq)) * a.mask_sq + q_base) * (size_t)a.mask_sk + n_abs;
#pragma unroll
          for (int c = 0; c < 32; ++c) mbias[c] = (kv_valid && q_base + c < I.sq) ? __ldg(mp + (size_t)c * a.mask_sk) : (uint16_t)0;
        }
        ptx::mbar_wait(b_q_full, g & 1);                 // LSE / D staged
        ptx::mbar_wait(b_s_full, g & 1);
        ptx::tc_fence_after();
        float p[32];
        {
          uint32_t(&r0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&p[0]);
          ptx::tmem_ld_x32(tS + lane_sel + w * 32, r0);
          ptx::tmem_wait_ld();
        }
        if constexpr (kMirror) {                 // S is in registers and nothing is written back: S(it + 1) may overwrite it
          ptx::tc_fence_before();
          ptx::mbar_arrive(b_p_ready);
        }
        // causal, fused / dK-dV: my key n_abs is visible to query q iff q >= lim; mirrored: my row sees key k iff k <= -lim
        const bool diag = a.causal && (kMirror ? (q_base + 31 > I.j * kT + I.shift) : (I.j * kT + kT - 1 > q_base + I.shift));   // warp-uniform
        const int lim = kMirror ? -(n_abs + I.shift) : n_abs - I.shift;
        const bool ragged = kMirror ? (q_base + 32 > I.sk) : ragged_k;
        float cs = c2;
        if constexpr (kMask) {                             // s <- s * scale_log2 + bias * log2 e
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const uint16_t mb = mbias[c];
            const float bias = kBF16 ? __uint_as_float((uint32_t)mb << 16) : __half2float(*reinterpret_cast<const __half *>(&mb));
            p[c] = fmaf(p[c], c2, bias * kLog2e);
          }
          cs = 1.f;
        }
        const float2 c2v = make_float2(cs, cs);
#pragma unroll
        for (int c = 0; c < 32; c += 4) {
          const float4 nl = kMirror ? make_float4(nl_mine, nl_mine, nl_mine, nl_mine)
                                    : ld_shared_f32x4(stat + 4 * c);       // -lse2 of 4 query columns
          const float2 x = __ffma2_rn(make_float2(p[c], p[c + 1]), c2v, make_float2(nl.x, nl.y));
          const float2 y = __ffma2_rn(make_float2(p[c + 2], p[c + 3]), c2v, make_float2(nl.z, nl.w));
          p[c] = ptx::ex2(x.x); p[c + 1] = ptx::ex2(x.y);
          p[c + 2] = ptx::ex2(y.x); p[c + 3] = ptx::ex2(y.y);
        }
        if (diag || ragged) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            if constexpr (kMirror) {                       // column = key q_base + c
              if (q_base + c >= I.sk || (diag && q_base + c > -lim)) p[c] = 0.f;
            } else {
              if (!kv_valid || (diag && q_base + c < lim)) p[c] = 0.f;
            }
          }
        }
        if constexpr (!kMirror) {
          uint32_t pk[16];
#pragma unroll
          for (int c = 0; c < 16; ++c) pk[c] = pack2(p[2 * c], p[2 * c + 1], kBF16);
          if constexpr (kDropout) {
            uint32_t wk[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) wk[k] = keep << (7 - k);
#pragma unroll
            for (int c = 0; c < 16; ++c) pk[c] &= pair_mask(wk, c);
          }
          ptx::tmem_st_x16(tS + lane_sel + w * 32, pk);     // P^T of my 32 queries: 16 packed columns inside my own S^T columns
          ptx::tmem_wait_st();
          ptx::tc_fence_before();
          ptx::mbar_arrive(b_p_ready);
        }

        ptx::mbar_wait(b_dp_full, g & 1);
        // (4) and (5) of the previous step have read dS^T: dq_full is committed after both (modes 1, 2: q_empty after (4))
        if (g > 0) ptx::mbar_wait(kMode == 0 ? b_dq_full : b_q_empty, (g - 1) & 1);
        ptx::tc_fence_after();
        {
          const uint32_t rowp = sbase + W::kDS + nrow * 128;
          uint32_t dpb[2][16];
          ptx::tmem_ld_x16(tDP + lane_sel + w * 32, dpb[0]);
          ptx::tmem_ld_x16(tDP + lane_sel + w * 32 + 16, dpb[1]);
          ptx::tmem_wait_ld();
#pragma unroll
          for (int q2 = 0; q2 < 2; ++q2) {
            uint32_t(&dp)[16] = dpb[q2];
            uint32_t ds[8];
#pragma unroll
            for (int c = 0; c < 16; c += 4) {
              const int cq = q2 * 16 + c;
              float2 g01 = ptx::u2f2(dp[c], dp[c + 1]), g23 = ptx::u2f2(dp[c + 2], dp[c + 3]);
              if constexpr (kDropout) {
                const uint32_t kb = keep >> cq;
                g01.x = (kb & 1u) ? g01.x : 0.f;
                g01.y = (kb & 2u) ? g01.y : 0.f;
                g23.x = (kb & 4u) ? g23.x : 0.f;
                g23.y = (kb & 8u) ? g23.y : 0.f;
              }
              const float4 nd = kMirror ? make_float4(nd_mine, nd_mine, nd_mine, nd_mine)
                                        : ld_shared_f32x4(stat + 256 + 4 * cq);   // -D (1 - p) of 4 query columns
              const float2 e01 = __fmul2_rn(make_float2(p[cq], p[cq + 1]), __fadd2_rn(g01, make_float2(nd.x, nd.y)));
              const float2 e23 = __fmul2_rn(make_float2(p[cq + 2], p[cq + 3]), __fadd2_rn(g23, make_float2(nd.z, nd.w)));
              ds[c / 2] = pack2(e01.x, e01.y, kBF16);
              ds[c / 2 + 1] = pack2(e23.x, e23.y, kBF16);
            }
#pragma unroll
            for (int k = 0; k < 2; ++k)     // my 32 queries = 64 bytes = chunks 4w .. 4w+3 of the key's 128-byte row
              st_shared_v4(rowp + (((4 * w + 2 * q2 + k) ^ (nrow & 7)) << 4), ds[4 * k], ds[4 * k + 1], ds[4 * k + 2], ds[4 * k + 3]);
          }
        }
        ptx::fence_proxy_async_smem();
        ptx::tc_fence_before();
        ptx::mbar_arrive(b_ds_ready);
      }
      g0 += I.n_it;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}

// dq = softmax_scale / (1 - p) * dq_accum^T: [64-query tile][32-query block][d_r][32] -> dq[token][head][d].
template <bool kBF16>
__global__ void __launch_bounds__(256)
dq_convert_wide_kernel(const float *__restrict__ acc, void *__restrict__ dq, const int32_t *__restrict__ cu_q,
                       int sq, int h, int nq64, float scale, int d, int d_r) {
  __shared__ float tile[8][32][33];
  const int i = blockIdx.x, hq = blockIdx.y, bidx = blockIdx.z;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  size_t tok0 = (size_t)bidx * sq;
  if (cu_q) {
    tok0 = (size_t)cu_q[bidx];
    sq = cu_q[bidx + 1] - cu_q[bidx];
    if (i * kQ >= sq) return;
  }
  const float *src = acc + (((size_t)bidx * h + hq) * nq64 + i) * 2 * d_r * 32;
  const int n_pairs = 2 * (d_r / 32);                  // (32-query block, 32-row group of d)
  for (int pair = warp; pair < n_pairs; pair += 8) {   // warp-uniform
    const int c = pair & 1, dg = pair >> 1;
    const float *blk = src + ((size_t)c * d_r + dg * 32) * 32;
#pragma unroll 8
    for (int rr = 0; rr < 32; ++rr)                    // row = d index, lane = query: one 128-byte line per row
      tile[warp][rr][lane] = blk[rr * 32 + ((((lane >> 2) ^ (rr & 7)) << 2) | (lane & 3))];
    __syncwarp();
    const int row = i * kQ + c * 32 + lane;            // my query
    if (row < sq) {
      uint16_t *dst = static_cast<uint16_t *>(dq) + ((tok0 + row) * h + hq) * d + dg * 32;
#pragma unroll
      for (int v8 = 0; v8 < 4; ++v8) {
        if (dg * 32 + v8 * 8 >= d) break;
        uint4 o;
        o.x = pack2(tile[warp][v8 * 8 + 0][lane] * scale, tile[warp][v8 * 8 + 1][lane] * scale, kBF16);
        o.y = pack2(tile[warp][v8 * 8 + 2][lane] * scale, tile[warp][v8 * 8 + 3][lane] * scale, kBF16);
        o.z = pack2(tile[warp][v8 * 8 + 4][lane] * scale, tile[warp][v8 * 8 + 5][lane] * scale, kBF16);
        o.w = pack2(tile[warp][v8 * 8 + 6][lane] * scale, tile[warp][v8 * 8 + 7][lane] * scale, kBF16);
        *reinterpret_cast<uint4 *>(dst + v8 * 8) = o;
      }
    }
    __syncwarp();
  }
}

std::atomic<SchedSlot *> g_bwdw_sched_addr[kMaxCachedDevices];
std::atomic<unsigned> g_bwdw_ticket{0};

bool wide_mask_tma_ok(const Params &p) {
  return p.mask && p.mask_sk % 8 == 0 && (reinterpret_cast<uintptr_t>(p.mask) & 15) == 0;
}

template <bool kBF16, bool kDropout, bool kMask, int kMode>
const char *launch_wide_t(const Params &p, cudaStream_t stream) {
  constexpr bool kMirror = kMode == 2;
  CUtensorMap tq, tk, tv, tdo, tdk, tdv;
  const char *e;
  const uint64_t nb = p.cu_q ? 1 : (uint64_t)p.b;
  if (!kMirror) {
    if ((e = make_tmap_rows(&tq, p.q, kBF16, p.d, p.h, p.sq, nb, kQ))) return e;
    if ((e = make_tmap_rows(&tdo, p.dout, kBF16, p.d, p.h, p.sq, nb, kQ))) return e;
    if ((e = make_tmap_rows(&tk, p.k, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
    if ((e = make_tmap_rows(&tv, p.v, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
    if ((e = make_tmap_rows(&tdk, p.dk, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
    if ((e = make_tmap_rows(&tdv, p.dv, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
  } else {
    // mirrored: the streamed pair (64-row boxes) is K, V; the resident pair (128-row boxes) Q, dO; the output dQ
    if ((e = make_tmap_rows(&tq, p.k, kBF16, p.d, p.hk, p.sk, nb, kQ))) return e;
    if ((e = make_tmap_rows(&tdo, p.v, kBF16, p.d, p.hk, p.sk, nb, kQ))) return e;
    if ((e = make_tmap_rows(&tk, p.q, kBF16, p.d, p.h, p.sq, nb, kT))) return e;
    if ((e = make_tmap_rows(&tv, p.dout, kBF16, p.d, p.h, p.sq, nb, kT))) return e;
    if ((e = make_tmap_rows(&tdk, p.dq, kBF16, p.d, p.h, p.sq, nb, kT))) return e;
    tdv = tdk;
  }

  WideArgs a;
  a.dq_accum = p.dq_accum; a.lse = p.lse; a.softmax_d = p.softmax_d;
  a.cu_q = p.cu_q; a.cu_k = p.cu_k;
  a.b = p.b; a.sq = p.sq; a.sk = p.sk; a.sq_rounded = p.sq_rounded; a.h = p.h; a.hk = p.hk;
  a.nq64 = (p.sq + kQ - 1) / kQ;
  a.batch_base = p.batch_base;
  a.d = p.d;
  a.d_r = (p.d + 31) / 32 * 32;
  a.scale = p.scale; a.scale_log2 = p.scale * kLog2e; a.rp_dropout = p.rp_dropout;
  a.keep_prob = kDropout ? 1.f - p.p_dropout : 1.f;
  a.drop_threshold = p.drop_threshold;
  a.key0 = (uint32_t)p.seed ^ (uint32_t)(p.offset >> 32);
  a.key1 = (uint32_t)(p.seed >> 32);
  a.ctr3 = (uint32_t)p.offset;
  a.causal = p.causal ? 1 : 0;
  a.mask = p.mask; a.mask_b = (int)p.mask_b; a.mask_h = (int)p.mask_h; a.mask_sq = (int)p.mask_sq; a.mask_sk = (int)p.mask_sk;
  a.hk_kv = p.hk;
  CUtensorMap tbias = tq;    // (a valid map either way; only read when mask_tma is set)
  a.mask_tma = 0;
  if (kMask && kMode == 1 && wide_mask_tma_ok(p)) {
    if ((e = make_tmap_bias(&tbias, p.mask, kBF16, (uint64_t)p.mask_sk, (uint64_t)p.mask_sq, (uint64_t)p.mask_h, (uint64_t)p.mask_b,
                            kT, kQ)))
      return e;
    a.mask_tma = 1;
  }

  auto kern = tc_bwd_wide_kernel<kBF16, kDropout, kMask, kMode>;
  int dev = 0, sm_count = 0;
  if ((e = current_device(&dev, &sm_count))) return e;
  static std::atomic<uint64_t> smem_opt_in{0};
  if ((e = ensure_dynamic_smem(kern, (int)W::kBytes, dev, smem_opt_in))) return e;
  a.nkt = (p.sk + kT - 1) / kT;
  a.n_items = 2 * a.nkt * p.hk * p.b;
  if (kMirror) {             // items are (128-query tile, QUERY head, batch, dh); the kernel reads them from nkt / hk
    a.nkt = (p.sq + kT - 1) / kT;
    a.hk = p.h;
    a.n_items = 2 * a.nkt * p.h * p.b;
  }
  a.persistent = a.n_items > sm_count ? 1 : 0;
  a.sched = nullptr;
  int grid = a.n_items;
  if (a.persistent) {
    grid = sm_count;
    if ((e = take_sched_slot(g_bwdw_sched, dev, g_bwdw_sched_addr, g_bwdw_ticket, &a.sched))) return e;
  }
  // (mode 0: dq_accum was zeroed by the preprocess kernel: tc_bwd_scratch_floats covers this layout)
  kern<<<grid, kThreads, W::kBytes, stream>>>(tq, tk, tv, tdo, tdk, tdv, tbias, a);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return cudaGetErrorString(err);
  if (kMode != 0) return nullptr;            // the two-launch backward never touches dq_accum
  dim3 cgrid(a.nq64, p.h, p.b);
  dq_convert_wide_kernel<kBF16><<<cgrid, 256, 0, stream>>>(p.dq_accum, p.dq, p.cu_q, p.sq, p.h, a.nq64,
                                                           p.scale * p.rp_dropout, a.d, a.d_r);
  err = cudaGetLastError();
  return err == cudaSuccess ? nullptr : cudaGetErrorString(err);
}

// num_splits == 1 or attn_mask: dK / dV, then the mirrored dQ pass, no atomics
template <bool kBF16, bool kDropout, bool kMask>
const char *launch_wide_two(const Params &p, cudaStream_t stream) {
  const char *e = launch_wide_t<kBF16, kDropout, kMask, 1>(p, stream);
  return e ? e : launch_wide_t<kBF16, kDropout, kMask, 2>(p, stream);
}

}  // namespace

const char *launch_tc_bwd_wide(const Params &p, cudaStream_t stream) {
  const bool drop = p.p_dropout > 0.f;
#define FA_WIDE(BF, DR)                                                                         \
  (p.mask ? launch_wide_two<BF, DR, true>(p, stream)                                            \
          : p.deterministic ? launch_wide_two<BF, DR, false>(p, stream)                         \
                            : launch_wide_t<BF, DR, false, 0>(p, stream))
  if (p.is_bf16) return drop ? FA_WIDE(true, true) : FA_WIDE(true, false);
  return drop ? FA_WIDE(false, true) : FA_WIDE(false, false);
#undef FA_WIDE
}

}  // namespace fa
