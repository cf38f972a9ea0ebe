// fa_tc_bwd.cu -- FlashAttention-2 backward for sm_100a (tcgen05 + TMEM + TMA).
// Implements flash_attn_bwd / flash_attn_varlen_bwd (reference flash_attn.h:68-99, :101-134) for head sizes up to 128
// (compiled widths 64 and 128; fa_tc_bwd_wide.cu takes 136..256).  The fused kernel described here is kMode 0; num_splits == 1
// and attn_mask run as two launches of the same kernel without atomics (kMode 1: dK / dV only, kMode 2: dQ only with the
// operand roles mirrored -- see the comment at the kernel).
//
// One CTA owns one 128-key tile (K_j, V_j) of one (batch, kv head) and loops over
// the query tiles that see it (and over the query heads of the GQA group),
// keeping dK_j and dV_j in tensor memory for the whole loop.  Scores are
// computed TRANSPOSED, so that TMEM lane == key row and P^T / dS^T are directly
// the A operands of the dV / dK GEMMs:
//   (1) S^T  = K_j Q_i^T                  A,B smem K-major        -> TMEM [0,128)
//   (2) dP^T = V_j dO_i^T                 A,B smem K-major        -> TMEM [128,256)
//       compute warps: P^T = exp2(S^T c - LSE_i), dS^T = P^T o (dP^T - D_i)
//   (3) dV_j += P^T dO_i                  A TMEM, B smem MN-major -> TMEM [256,256+D)
//   (4) dK_j += dS^T Q_i                  A smem K-major (written by threads), B MN-major
//                                                                 -> TMEM [384,384+D)
//   (5) dQ_i  = dS K_j                    A smem MN-major (same buffer), B MN-major
//                                                                 -> TMEM [128,128+D) (reuses dP^T)
// dQ_i tiles are added into the fp32 scratch `dq_accum` with bulk reduce-adds
// (cp.reduce.async.bulk .add.f32) of 16 KB blocks; a small kernel converts the
// scratch to dq.  Sixteen warps:
//   warps 0-7   compute (lane = key row; warps 0-3 own query columns 0-63, 4-7 own 64-127)
//   warp  8     tcgen05.mma issue, TMEM alloc/free     warp 9   TMA producer (+ LSE/D staging)
//   warps 12-15 dQ drain (TMEM -> smem -> bulk reduce-add)
#include "fa_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"
#include "fa_bwd_dev.cuh"
#include "fa_trace.cuh"

namespace fa {

namespace {

#ifndef FA_BWD_SYNC_GROUPS
#define FA_BWD_SYNC_GROUPS 2
#endif
#ifndef FA_BWD_EMU_EVERY
#define FA_BWD_EMU_EVERY 0
#endif
#ifndef FA_BWD_DRAIN_REGS
#define FA_BWD_DRAIN_REGS 144
#endif
constexpr int kBwdSyncGroups = FA_BWD_SYNC_GROUPS;   // dropout: scheduling block = this many groups of 8 exponentials + 5 Philox rounds
// Every n-th group of four exponentials computes one pair on the FMA pipe instead of MUFU (0 = never).
constexpr int kBwdEmuEvery = FA_BWD_EMU_EVERY;
constexpr int kPhiloxSteps = 4 * FA_PHILOX_ROUNDS;   // round steps of an iteration's four calls, spread over its 8 groups
constexpr int kT = 128;            // tile edge: keys per CTA tile, queries per inner tile
constexpr int kThreads = 512;
constexpr int kMmaWarp = 8, kLoadWarp = 9, kDrainWarp0 = 12;
constexpr float kLog2e = 1.4426950408889634f;

template <int D>
struct Cfg {
  static constexpr uint32_t kTile = kT * D * 2;       // one 128 x D 16-bit tile
  static constexpr uint32_t kSub = kT * 128;          // one [128 x 64] swizzled sub-tile
  static constexpr uint32_t kK = 0;
  static constexpr uint32_t kV = kK + kTile;
  static constexpr uint32_t kQ = kV + kTile;          // 2 stages
  static constexpr uint32_t kDO = kQ + 2 * kTile;
  static constexpr uint32_t kDS = kDO + kTile;        // dS^T: 128 x 128 16-bit = 32 KB
  static constexpr uint32_t kDQ = kDS + 2 * kSub;     // 2 x 16 KB fp32 staging blocks
  static constexpr uint32_t kStat = kDQ + 2 * 16384;  // 2 stages x {lse2[128], D[128]} fp32
  static constexpr uint32_t kBar = kStat + 2 * 1024;
  static constexpr uint32_t kNumBars = 22;
  static constexpr uint32_t kDesc = kBar + 256;        // 6 x 128 B tensor-map staging (packed inputs)
  static constexpr uint32_t kBytes = kDesc + 6 * 128;
  static_assert(kBytes <= 232448, "over the 227 KB a CTA may ask for");
  static_assert(kBytes > 116 * 1024, "one CTA per SM is what makes the per-SM tensor-map slots private");
};

__device__ SchedSlot g_bwd_sched[kSchedSlots];       // item counters of the persistent launches (fa_common.cuh)
__device__ ptx::TensorMapSlot g_bwd_slots[ptx::kMaxSm][2][6];   // two alternating sets (packed inputs)

struct BwdArgs {
  void *dk, *dv;
  float *dq_accum;
  const float *lse, *softmax_d;
  const int32_t *cu_q, *cu_k;
  int b, sq, sk, sq_rounded, h, hk, nqt, batch_base;
  const void *mask;      // additive bias [mask_b, mask_h, mask_sq, mask_sk] in the input dtype, or null; packed inputs:
  int mask_b, mask_h;    // mask_sq / mask_sk = max_seqlen_q / max_seqlen_k, entry (row, col) counted from the sequence's start
  int mask_sq, mask_sk;
  int d, nblk;           // real head_size (<= D, multiple of 8; the rest is TMA zero fill); 32-column blocks of dq_accum   // sq/sk/nqt: dense values or packed maxima
  float scale, scale_log2, rp_dropout;
  float keep_prob;       // 1 - p_dropout.  dS is computed as P o (keep o dP - D (1 - p)) = (1 - p) dS; the factor 1 / (1 - p)
                         // rides on the final scales of dQ and dK (one multiply per output element, not per score)
  uint32_t drop_threshold, key0, key1, ctr3;
  int causal;
  int mask_tma;          // kMode 1 with attn_mask: bias tiles arrive by TMA (tmap_bias)
  int hk_kv;             // K/V heads (mode 2 overwrites hk with the query head count: its items are per query head)
  int nkt, n_items;      // key tiles per sequence (dense: of sk; packed: of the longest); items = nkt * hk * b
  int persistent;        // 1: CTAs pull items from *sched until none are left; 0: item = blockIdx.x
  SchedSlot *sched;      // global item counter of this launch (self-resetting pool slot, fa_common.cuh)
};

using namespace bwd_dev;

}  // namespace

namespace {

struct BwdBars {
  uint32_t kv_full, q_full /*[2]*/, q_empty /*[2]*/, do_full, do_empty, s_full, dp_full, p_ready, ds_ready,
      dq_full, dq_drained, acc_done, epi_tmem_free, epi_smem_free, sched_full /*[2]*/, sched_empty /*[2]*/,
      bias_full, bias_empty;
};

// kMode: 0 = the fused kernel described at the top (dQ by reduce-adds in arrival order).
// Deterministic dQ (num_splits == 1, reference flash_attn.h:94) runs as TWO launches of this kernel without any atomic:
//   1 = dK / dV only: the fused kernel minus GEMM (5) and the dQ drain;
//   2 = dQ only, MIRRORED: one item = one 128-QUERY tile of one (batch, query head), resident operands Q and dO
//       (in the K and V slots), streamed operands K and V (in the Q and dO slots): the same GEMMs then give
//       (1) S = Q K^T, (2) dP = dO V^T with TMEM lane = QUERY row, and (4) accumulates dQ = dS K in the dK columns.
//       LSE / D are per LANE (two scalars per item instead of staged vectors), the causal mask and the dropout draw
//       are the row-wise ones of the forward (no bit transposes), nothing leaves through dq_accum.
// 7 GEMM units per tile pair instead of 5 and twice the exponentials, but no serialised turns: round 1's ordered
// reduce-adds (a turn counter per query tile) ran at 423 / 565 TFLOP/s causal / full because the link between two
// consecutive adders of a tile (adds performed, fence, release, acquire poll: ~7 us) sat on every step; an upward
// walk with turn order = arrival order was measured too: 305 / 513 (profiles/r02c_ordered_dq.log).
template <int D, bool kBF16, bool kDropout, bool kMask, int kMode>
__global__ void __launch_bounds__(kThreads, 1)
tc_bwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
              const __grid_constant__ CUtensorMap tmap_v, const __grid_constant__ CUtensorMap tmap_do,
              const __grid_constant__ CUtensorMap tmap_dk, const __grid_constant__ CUtensorMap tmap_dv,
              const __grid_constant__ CUtensorMap tmap_bias, const BwdArgs a) {
  using C = Cfg<D>;
  // attn_mask in the dK / dV pass (kMode 1): the pass has no dQ to drain, so the 32 KB of dQ staging are free during
  // an item's loop and hold the step's [128 query x 128 key] bias tile, brought by TMA one step ahead (a lane owns a
  // key COLUMN of the bias here: read from global memory that is 64 two-byte loads per lane and step, latency exposed)
  constexpr bool kBiasTma = kMask && kMode == 1;
  constexpr bool kMirror = kMode == 2;
  // register split of the 16 warps (2048 per lane in total): without dropout the drain warps
  // hold a whole dQ tile; with dropout the compute warps need the room for the keep-bit exchange
  constexpr bool kEarlyDrain = true;
  // drain / compute registers per thread; 4 x 32 (producers) + 4 x drain + 8 x compute = 2048
  constexpr int kDrainRegs = D == 64 ? 128 : ((kDropout || kMask) ? FA_BWD_DRAIN_REGS : 208);   // bias: 64 prefetched words
  constexpr int kComputeRegs = (2048 - 128 - 4 * kDrainRegs) / 8;
  static_assert(kComputeRegs % 8 == 0 && kDrainRegs % 8 == 0, "setmaxnreg takes multiples of 8");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t sbase = ptx::smem_u32(smem_raw);
  if (sbase & 1023u) __trap();   // the swizzled tiles need a 1024-byte aligned base
  // warp index through a shuffle: the role branches are then provably warp-uniform, and tcgen05.mma under
  // elect.sync is issued without a per-instruction election loop (see fa_tc_fwd.cu)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  // ---- work items -----------------------------------------------------------------------
  // One item = one 128-key tile of one (batch, kv head): dK_j / dV_j stay in tensor memory while the item's
  // loop walks the group's query heads and the query tiles that see the key tile.  Linear order: key tiles
  // fastest (causal: the heaviest first).  Dense inputs: one persistent CTA per SM pulls items from a global
  // counter, and its TMA / MMA / drain warps run ahead into the next item while the compute warps write dK
  // and dV out (a CTA's start -- DRAM latency of the first loads -- and its 64 KB epilogue on a ~26 B/clk
  // store port were ~8000 of ~70000 cycles per item, traced).  Packed inputs run persistent too (tensor maps re-bounded per sequence).
  struct Item {
    int j, hk, bidx, sq, sk, qrow0, krow0, bcoord, shift, i_min, n_i, n_it;
    size_t qtok0, ktok0;
    bool valid;
  };
  const int group_real = a.h / a.hk_kv;
  const int group = kMirror ? 1 : group_real;          // mirrored: `hk` below IS the query head, no loop over a group
  auto decode = [&](int L) {
    Item I;
    I.j = L % a.nkt; I.hk = (L / a.nkt) % a.hk; I.bidx = L / (a.nkt * a.hk);
    if constexpr (kMirror) {                           // a.nkt / a.hk hold the query tile and query head counts
      I.j = a.causal ? a.nkt - 1 - I.j : I.j;          // causal: the last query tile sees the most keys: heaviest first
    }
    I.sq = a.sq; I.sk = a.sk; I.qrow0 = 0; I.krow0 = 0; I.bcoord = I.bidx;   // TMA row origin / batch coordinate
    I.qtok0 = (size_t)I.bidx * a.sq; I.ktok0 = (size_t)I.bidx * a.sk;        // first token of this sequence
    if (a.cu_q) {
      I.qrow0 = a.cu_q[I.bidx]; I.krow0 = a.cu_k[I.bidx];
      I.sq = a.cu_q[I.bidx + 1] - I.qrow0; I.sk = a.cu_k[I.bidx + 1] - I.krow0;
      I.qtok0 = (size_t)I.qrow0; I.ktok0 = (size_t)I.krow0; I.bcoord = 0;
    }
    I.shift = I.sk - I.sq;                       // causal: key n visible to row r iff n <= r + shift
    if constexpr (kMirror) {
      // `j` = my QUERY tile, the loop walks the key tiles 0 .. n_it - 1 that its rows see
      I.valid = I.j * kT < I.sq;
      int nk = (I.sk + kT - 1) / kT;
      if (a.causal) {
        const int last_key = I.j * kT + kT - 1 + I.shift;      // last key the tile's last row sees
        nk = last_key < 0 ? 0 : min(nk, last_key / kT + 1);
      }
      I.i_min = 0; I.n_i = nk; I.n_it = nk;
      return I;
    }
    I.valid = I.j * kT < I.sk;                   // packed: this sequence may have no such key tile
    const int nqt = (I.sq + kT - 1) / kT;        // query tiles of this sequence
    I.i_min = 0;
    if (a.causal) {
      const int first_row = I.j * kT - I.shift;  // first query row that sees key j*128
      I.i_min = first_row <= 0 ? 0 : first_row / kT;
    }
    I.n_i = max(0, nqt - I.i_min);               // query tiles per head
    I.n_it = I.n_i * group;                      // inner iterations: (head in group, query tile)
    return I;
  };
  if (!a.persistent && !decode((int)blockIdx.x).valid) return;
  // Query tile of inner iteration `it` (mirrored: key tile).
  constexpr bool descending = false;
#define FA_ITEM_LOCALS(I)                                                                                      \
  const int j = (I).j, hk = (I).hk, bidx = (I).bidx, sq = (I).sq, sk = (I).sk, qrow0 = (I).qrow0,              \
            krow0 = (I).krow0, bcoord = (I).bcoord, shift = (I).shift, i_min = (I).i_min, n_i = (I).n_i,      \
            n_it = (I).n_it;                                                                                   \
  const size_t qtok0 = (I).qtok0, ktok0 = (I).ktok0;                                                           \
  auto q_tile_of = [&](int it_) { return descending ? i_min + n_i - 1 - it_ % n_i : i_min + it_ % n_i; };      \
  (void)j; (void)hk; (void)bidx; (void)sq; (void)sk; (void)qrow0; (void)krow0; (void)bcoord; (void)shift;      \
  (void)i_min; (void)qtok0; (void)ktok0; (void)q_tile_of

  // ---- barriers, tensor memory --------------------------------------------------------
  const uint32_t bar0 = sbase + C::kBar;
  BwdBars B;
  B.kv_full = bar0;          B.q_full = bar0 + 8;      B.q_empty = bar0 + 24;   B.do_full = bar0 + 40;
  B.do_empty = bar0 + 48;    B.s_full = bar0 + 56;     B.dp_full = bar0 + 64;   B.p_ready = bar0 + 72;
  B.ds_ready = bar0 + 80;    B.dq_full = bar0 + 88;    B.dq_drained = bar0 + 96; B.acc_done = bar0 + 104;
  B.epi_tmem_free = bar0 + 112; B.epi_smem_free = bar0 + 120; B.sched_full = bar0 + 128; B.sched_empty = bar0 + 144;
  B.bias_full = bar0 + 160;  B.bias_empty = bar0 + 168;
  static_assert(C::kNumBars >= 22 && C::kNumBars * 8 + 32 <= 256, "barrier block too small");
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem_raw + C::kBar + C::kNumBars * 8);
  volatile int *sched_item = reinterpret_cast<volatile int *>(tmem_slot + 2);   // [2]

  if (threadIdx.x == 0) {
    ptx::mbar_init(B.kv_full, 1);
    ptx::mbar_init(B.q_full, 2);      ptx::mbar_init(B.q_full + 8, 2);    // TMA expect_tx arrive + stats arrive
    ptx::mbar_init(B.q_empty, 1);     ptx::mbar_init(B.q_empty + 8, 1);
    ptx::mbar_init(B.do_full, 1);     ptx::mbar_init(B.do_empty, 1);
    ptx::mbar_init(B.s_full, 1);      ptx::mbar_init(B.dp_full, 1);
    ptx::mbar_init(B.p_ready, kMirror ? 384 : 256);   ptx::mbar_init(B.ds_ready, kMirror ? 384 : 256);   // mirrored: 12 warps
    ptx::mbar_init(B.dq_full, 1);     ptx::mbar_init(B.dq_drained, 128);
    ptx::mbar_init(B.acc_done, 1);
    ptx::mbar_init(B.epi_tmem_free, 128);  ptx::mbar_init(B.epi_smem_free, 1);   // (epi_smem_free: bias tiles only)
    ptx::mbar_init(B.bias_full, 1);        ptx::mbar_init(B.bias_empty, 256);
    for (int t = 0; t < 2; ++t) {
      ptx::mbar_init(B.sched_full + 8 * t, 1);
      ptx::mbar_init(B.sched_empty + 8 * t, 1 + 128 + 256);    // MMA lane, drain warps, compute warps
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
  const uint32_t tS = tmem, tDP = tmem + 128, tDV = tmem + 256, tDK = tmem + 384;

#ifdef FA_TRACE
  const bool tr_cta = blockIdx.x == gridDim.x / 2;
#endif
  // Consumers of the item queue (MMA lane, drain and compute threads): item number n of this CTA, or -1.
  auto next_item = [&](int n) {
    const int slot = n & 1;
    ptx::mbar_wait(B.sched_full + 8 * slot, (n >> 1) & 1);
    const int L = sched_item[slot];
    ptx::mbar_arrive(B.sched_empty + 8 * slot);
    return L;
  };
  // ---- mirrored dQ pass (mode 2): one item's score work on the NC key columns [c0, c0 + NC) of every step.
  // TWELVE warps share a step's 128 columns 48 / 48 / 32: the eight compute warps and the four drain warps, which have
  // nothing to drain in this mode (under ncu the pass ran at 59 % tensor-pipe activity on eight: bound by exactly this work).
  // A lane owns one QUERY row: LSE / D are two scalars, the causal mask and the Philox calls are the forward's row-wise ones.
  auto mirror_item = [&](auto nc_tag, const int c0, const Item &I, const int g0) {
    constexpr int NC = decltype(nc_tag)::value;
    static_assert(NC == 48 || NC == 32, "column split of the mirrored pass");
    FA_ITEM_LOCALS(I);
    const int n = (warp & 3) * 32 + lane;              // query row inside the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    const float c2 = a.scale_log2;
    uint32_t add7 = 0;
    bool t_msb = false;
    if constexpr (kDropout) {
      add7 = (0x80u - (a.drop_threshold & 0x7fu)) * 0x01010101u;
      t_msb = a.drop_threshold >= 128u;
    }
    const int n_abs = j * kT + n;
    const bool row_ok = n_abs < sq;
    float nl_mine = -INFINITY, nd_mine = 0.f;          // -LSE log2 e (-inf past the end => P = 0), -D (1 - p)
    if (row_ok) {
      nl_mine = -a.lse[((size_t)bidx * a.h + hk) * a.sq + n_abs] * kLog2e;
      nd_mine = -a.softmax_d[((size_t)bidx * a.h + hk) * a.sq_rounded + n_abs] * a.keep_prob;
    }
    const uint32_t bh = (uint32_t)((a.batch_base + bidx) * a.h + hk);
    for (int it = 0; it < n_it; ++it) {
      const int i = it, g = g0 + it;
      const int k_base = i * kT + c0;                  // my first key column
      uint32_t keep[2] = {0xffffffffu, 0xffffffffu};   // bit t of word t / 32: keep(my row, key k_base + t)
      if constexpr (kDropout) {
        keep[0] = 0; keep[1] = 0;
#pragma unroll
        for (int blk = 0; blk < NC / 16; ++blk) {
          const fa_philox4_t rnd = fa_philox4x32((uint32_t)n_abs, (uint32_t)((k_base + 16 * blk) >> 4), bh, a.ctr3, a.key0, a.key1);
          keep[blk / 2] |= keep_bits16(rnd, add7, t_msb) << (16 * (blk & 1));
        }
      }
      uint16_t mbias[kMask ? NC : 1];
      if constexpr (kMask) {                           // my row's NC bias values: contiguous bytes
        const uint16_t *mp = static_cast<const uint16_t *>(a.mask) +
            (((size_t)(a.mask_b == 1 ? 0 : bidx) * a.mask_h + (a.mask_h == 1 ? 0 : hk)) * a.mask_sq + n_abs) * (size_t)a.mask_sk + k_base;
        const bool fast = row_ok && k_base + NC <= sk && a.mask_sk % 8 == 0 && (reinterpret_cast<uintptr_t>(a.mask) & 15) == 0;
        if (fast) {
#pragma unroll
          for (int v8 = 0; v8 < NC / 8; ++v8) {
            const uint4 x = __ldg(reinterpret_cast<const uint4 *>(mp) + v8);
            const uint32_t w4[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) mbias[8 * v8 + e] = (uint16_t)(w4[e / 2] >> (16 * (e & 1)));
          }
        } else {
#pragma unroll
          for (int c = 0; c < NC; ++c) mbias[c] = (row_ok && k_base + c < sk) ? __ldg(mp + c) : (uint16_t)0;
        }
      }
      ptx::mbar_wait(B.s_full, g & 1);
      ptx::tc_fence_after();
      float p[NC];
      {
        uint32_t(&r0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&p[0]);
        ptx::tmem_ld_x32(tS + lane_sel + c0, r0);
        if constexpr (NC == 48) {
          uint32_t(&r1)[16] = *reinterpret_cast<uint32_t(*)[16]>(&p[32]);
          ptx::tmem_ld_x16(tS + lane_sel + c0 + 32, r1);
        }
        ptx::tmem_wait_ld();
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(B.p_ready);                     // S is in registers: S(it + 1) may overwrite it
      float cs = c2;
      if constexpr (kMask) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
          const uint16_t mb = mbias[c];
          const float bias = kBF16 ? __uint_as_float((uint32_t)mb << 16) : __half2float(*reinterpret_cast<const __half *>(&mb));
          p[c] = fmaf(p[c], c2, bias * kLog2e);
        }
        cs = 1.f;
      }
      const float2 c2v = make_float2(cs, cs), nlv = make_float2(nl_mine, nl_mine);
#pragma unroll
      for (int c = 0; c < NC; c += 2) {
        const float2 x = __ffma2_rn(make_float2(p[c], p[c + 1]), c2v, nlv);
        p[c] = ptx::ex2(x.x); p[c + 1] = ptx::ex2(x.y);
      }
      const bool diag = a.causal && (k_base + NC - 1 > j * kT + shift);   // warp-uniform: some key past the tile's first row's limit
      if (diag || k_base + NC > sk) {
#pragma unroll
        for (int c = 0; c < NC; ++c)
          if (k_base + c >= sk || (diag && k_base + c > n_abs + shift)) p[c] = 0.f;
      }
      ptx::mbar_wait(B.dp_full, g & 1);
      if (g > 0) ptx::mbar_wait(B.q_empty + 8 * ((g - 1) & 1), ((g - 1) >> 1) & 1);   // (4) of it - 1 has read dS
      ptx::tc_fence_after();
      {
        const float2 ndv = make_float2(nd_mine, nd_mine);
        uint32_t dpb[2][16];
        ptx::tmem_ld_x16(tDP + lane_sel + c0, dpb[0]);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int q4 = 0; q4 < NC / 16; ++q4) {
          uint32_t(&dp)[16] = dpb[q4 & 1];
          if (q4 + 1 < NC / 16) ptx::tmem_ld_x16(tDP + lane_sel + c0 + (q4 + 1) * 16, dpb[(q4 + 1) & 1]);
          uint32_t ds[8];
#pragma unroll
          for (int c = 0; c < 16; c += 2) {
            const int cq = q4 * 16 + c;
            float2 gg = ptx::u2f2(dp[c], dp[c + 1]);
            if constexpr (kDropout) {
              const uint32_t kb = keep[cq / 32] >> (cq % 32);
              gg.x = (kb & 1u) ? gg.x : 0.f;
              gg.y = (kb & 2u) ? gg.y : 0.f;
            }
            const float2 e = __fmul2_rn(make_float2(p[cq], p[cq + 1]), __fadd2_rn(gg, ndv));
            ds[c / 2] = pack2(e.x, e.y, kBF16);
          }
#pragma unroll
          for (int k = 0; k < 2; ++k) {                // logical 16-byte chunk (8 keys) of my row of dS [128 q][128 keys]
            const int idx = (c0 + q4 * 16) / 8 + k;
            st_shared_v4(sbase + C::kDS + (idx / 8) * C::kSub + n * 128 + (((idx % 8) ^ (n & 7)) << 4),
                         ds[4 * k], ds[4 * k + 1], ds[4 * k + 2], ds[4 * k + 3]);
          }
          if (q4 + 1 < NC / 16) ptx::tmem_wait_ld();
        }
      }
      ptx::fence_proxy_async_smem();
      ptx::tc_fence_before();
      ptx::mbar_arrive(B.ds_ready);
    }
  };

  if (warp >= 8 && warp < 12) {
    // =========================== producer warpgroup ======================================
    ptx::setmaxnreg_dec<32>();
    if (warp == kLoadWarp) {
      // ---- TMA producer; all 32 lanes stage LSE / D -----------------------------------------
      const CUtensorMap *mq = &tmap_q, *mk = &tmap_k, *mv = &tmap_v, *mdo = &tmap_do;
      int mapped_b = -1, set = 0, nv = 0;              // packed: sequence the maps are bounded to; valid items so far
      int g0 = 0;                                      // iterations of the items before this one (ring stages, phases)
      for (int n = 0;; ++n) {
      // ---- scheduler: lane 0 takes the next item and publishes it to the other roles ----
      int L = 0;
      if (lane == 0) {
        const int slot = n & 1;
        if (n >= 2) ptx::mbar_wait(B.sched_empty + 8 * slot, ((n >> 1) - 1) & 1);
        L = a.persistent ? sched_next_item(a.sched, a.n_items) : (n == 0 ? (int)blockIdx.x : -1);
        sched_item[slot] = L;
        ptx::mbar_arrive(B.sched_full + 8 * slot);
      }
      L = __shfl_sync(0xffffffffu, L, 0);
      if (L < 0) break;
      const Item I = decode(L);
      if (!I.valid) continue;                          // packed: this sequence has no such key tile
      FA_ITEM_LOCALS(I);
      if (nv >= 1) ptx::mbar_wait(B.acc_done, (nv - 1) & 1);   // the previous item's MMAs have read K and V
      if (a.cu_q && bidx != mapped_b) {
        // packed inputs: rows past this sequence's end must read as zero, not as the next sequence.  Every load
        // issued so far has been consumed (acc_done above), so the other slot set may be rewritten.
        const uint32_t sm = ptx::smid();
        if (sm >= (uint32_t)ptx::kMaxSm) __trap();
        const uint32_t d0 = sbase + C::kDesc;
        uint8_t *gen0 = smem_raw + C::kDesc;
        mq = ptx::patch_rows(tmap_q, &g_bwd_slots[sm][set][0], d0, gen0, (uint32_t)(qrow0 + sq), lane);
        mdo = ptx::patch_rows(tmap_do, &g_bwd_slots[sm][set][1], d0 + 128, gen0 + 128, (uint32_t)(qrow0 + sq), lane);
        mk = ptx::patch_rows(tmap_k, &g_bwd_slots[sm][set][2], d0 + 256, gen0 + 256, (uint32_t)(krow0 + sk), lane);
        mv = ptx::patch_rows(tmap_v, &g_bwd_slots[sm][set][3], d0 + 384, gen0 + 384, (uint32_t)(krow0 + sk), lane);
        mapped_b = bidx;
        set ^= 1;
      }
      // resident pair (K and V slots) and streamed pair (Q and dO slots): K, V | Q, dO -- mirrored: Q, dO | K, V
      const CUtensorMap *res_a = kMirror ? mq : mk, *res_b = kMirror ? mdo : mv;
      const CUtensorMap *str_a = kMirror ? mk : mq, *str_b = kMirror ? mv : mdo;
      const int res_row0 = kMirror ? qrow0 : krow0, str_row0 = kMirror ? krow0 : qrow0;
      if (lane == 0) {
        ptx::mbar_arrive_expect_tx(B.kv_full, 2 * C::kTile);
        for (int c = 0; c < D / 64; ++c) {
          ptx::tma_load_4d(sbase + C::kK + c * C::kSub, res_a, B.kv_full, c * 64, hk, res_row0 + j * kT, bcoord);
          ptx::tma_load_4d(sbase + C::kV + c * C::kSub, res_b, B.kv_full, c * 64, hk, res_row0 + j * kT, bcoord);
        }
      }
      for (int it = 0; it < n_it; ++it) {
        const int hq = hk * group + it / n_i, i = q_tile_of(it);
        const int str_head = kMirror ? hk / group_real : hq;
        const int g = g0 + it, s = g & 1;
        if constexpr (kBiasTma) {
          if (a.mask_tma) {
            if (g >= 1) ptx::mbar_wait(B.bias_empty, (g - 1) & 1);                  // the compute warps have read tile g - 1
            if (it == 0 && nv >= 1) ptx::mbar_wait(B.epi_smem_free, (nv - 1) & 1);  // the last item's dK / dV have left the staging
            if (lane == 0) {
              ptx::mbar_arrive_expect_tx(B.bias_full, kT * kT * 2);
              ptx::tma_load_4d(sbase + C::kDQ, &tmap_bias, B.bias_full, j * kT, i * kT, a.mask_h == 1 ? 0 : hq,
                               a.mask_b == 1 ? 0 : bidx);
            }
          }
        }
        if (g >= 2) ptx::mbar_wait(B.q_empty + 8 * s, ((g >> 1) - 1) & 1);
        if (lane == 0) {
          ptx::mbar_arrive_expect_tx(B.q_full + 8 * s, C::kTile);
          for (int c = 0; c < D / 64; ++c)
            ptx::tma_load_4d(sbase + C::kQ + s * C::kTile + c * C::kSub, str_a, B.q_full + 8 * s,
                             c * 64, str_head, str_row0 + i * kT, bcoord);
        }
        if constexpr (kMirror) {
          if (lane == 0) ptx::mbar_arrive(B.q_full + 8 * s);     // (LSE / D are per lane here: nothing to stage)
        } else {
          // staged NEGATED: -LSE * log2(e) (-inf past the sequence end => P = 0 there), -D = -rowsum(dO o O)
          const float *lse = a.lse + ((size_t)bidx * a.h + hq) * a.sq;
          const float *dd = a.softmax_d + ((size_t)bidx * a.h + hq) * a.sq_rounded;
          const uint32_t st = sbase + C::kStat + s * 1024;
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const int r = t * 32 + lane, row = i * kT + r;
            const float l = row < sq ? lse[row] : INFINITY;
            const float dv = row < sq ? dd[row] : 0.f;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(st + 4 * r), "f"(-l * kLog2e) : "memory");
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(st + 512 + 4 * r), "f"(-dv * a.keep_prob) : "memory");
          }
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(B.q_full + 8 * s);
        }
        if (g >= 1) ptx::mbar_wait(B.do_empty, (g - 1) & 1);
        if (lane == 0) {
          ptx::mbar_arrive_expect_tx(B.do_full, C::kTile);
          for (int c = 0; c < D / 64; ++c)
            ptx::tma_load_4d(sbase + C::kDO + c * C::kSub, str_b, B.do_full, c * 64, str_head, str_row0 + i * kT, bcoord);
        }
      }
      if constexpr (kBiasTma) {      // (every phase of epi_smem_free is waited for, in order: an item without steps has no first bias load to do it)
        if (a.mask_tma && n_it == 0 && nv >= 1) ptx::mbar_wait(B.epi_smem_free, (nv - 1) & 1);
      }
      g0 += n_it;
      ++nv;
      }   // items
    } else if (warp == kMmaWarp && ptx::elect_one()) {
      // ---- MMA issuer -----------------------------------------------------------------------
      constexpr uint32_t fmt = kBF16 ? ptx::kFmtBF16 : ptx::kFmtF16;
      constexpr uint32_t id_kk = ptx::make_idesc(fmt, ptx::kMajorK, ptx::kMajorK, kT, kT);     // (1) (2)
      constexpr uint32_t id_k_mn = ptx::make_idesc(fmt, ptx::kMajorK, ptx::kMajorMN, kT, D);   // (3) (4)
      constexpr uint32_t id_mn_mn = ptx::make_idesc(fmt, ptx::kMajorMN, ptx::kMajorMN, kT, D); // (5)
      constexpr uint64_t hi_k = ptx::make_smem_desc_hi(16, 1024);
      constexpr uint64_t hi_mn = ptx::make_smem_desc_hi(C::kSub, 1024);
      // D(tmem) = A(K-major smem) * B(K-major smem), contraction over d; a_off/b_off: bytes from sbase
      auto gemm_kk = [&](uint32_t d_tmem, uint32_t a16, uint32_t a_off, uint32_t b16, uint32_t b_off) {
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          const uint32_t off = (ks / 4) * C::kSub + (ks % 4) * 32;
          ptx::mma_ss(d_tmem, ptx::smem_desc_at(hi_k, a16, a_off + off), ptx::smem_desc_at(hi_k, b16, b_off + off),
                      id_kk, ks > 0);
        }
      };
      FA_TRACE_DECL(tr, 0, tr_cta);
      int g0 = 0, nv = 0;
      for (int n = 0;; ++n) {
      const int L = next_item(n);
      if (L < 0) break;
      const Item I = decode(L);
      if (!I.valid) continue;
      FA_ITEM_LOCALS(I);
      ptx::mbar_wait(B.kv_full, nv & 1);
      ptx::tc_fence_after();
      // An item without steps (mirrored pass: a query tile that sees no key) commits acc_done at once.  The drain warps wait
      // for acc_done by PARITY: were this warp to run two items ahead of them, they would wait for the wrong phase (seen as a
      // hang on a packed causal batch with sk << sq: runs of such items).  So an empty item first waits for the previous
      // item's epilogue, which a non-empty one does before its first accumulating MMA anyway.
      if (n_it == 0 && nv >= 1) ptx::mbar_wait(B.epi_tmem_free, (nv - 1) & 1);
      if (n_it > 0) {
        ptx::mbar_wait(B.q_full + 8 * (g0 & 1), (g0 >> 1) & 1);
        ptx::tc_fence_after();
        gemm_kk(tS, sbase >> 4, C::kK, (sbase >> 4) + (g0 & 1) * (C::kTile >> 4), C::kQ);   // (1) S^T(0)
        ptx::mma_commit(B.s_full);
      }
      for (int it = 0; it < n_it; ++it) {
        const int g = g0 + it, s = g & 1;
        // descriptors are rebuilt from this laundered base every iteration (see smem_desc_at)
        const uint32_t b16 = ptx::opaque(sbase >> 4);
        const uint32_t q16 = b16 + s * (C::kTile >> 4), qn16 = b16 + (s ^ 1) * (C::kTile >> 4);
        FA_TRACE_EV(tr, 1);
        ptx::mbar_wait(B.do_full, g & 1);
        FA_TRACE_EV(tr, 2);
        if (kMode == 0 && g > 0) ptx::mbar_wait(B.dq_drained, (g - 1) & 1);   // dP^T shares TMEM with dQ
        ptx::tc_fence_after();
        FA_TRACE_EV(tr, 3);
        gemm_kk(tDP, b16, C::kV, b16, C::kDO);                               // (2) dP^T(it)
        ptx::mma_commit(B.dp_full);
        if constexpr (kMirror) ptx::mma_commit(B.do_empty);                  // no (3) here: the dO slot is free after (2)
        ptx::mbar_wait(B.p_ready, g & 1);                                    // (mirrored: S has been read, nothing written)
        // the first dV / dK MMAs of an item overwrite the accumulators: the previous item's epilogue must have read them
        if (it == 0 && nv >= 1) ptx::mbar_wait(B.epi_tmem_free, (nv - 1) & 1);
        ptx::tc_fence_after();
        FA_TRACE_EV(tr, 4);
        if constexpr (!kMirror) {
#pragma unroll
          for (int ks = 0; ks < kT / 16; ++ks)                             // (3) dV += P^T dO
            ptx::mma_ts(tDV, tS + (ks / 4) * 64 + (ks % 4) * 8,
                        ptx::smem_desc_at(hi_mn, b16, C::kDO + ks * 2048), id_k_mn, it > 0 || ks > 0);
          ptx::mma_commit(B.do_empty);
        }
        if (it + 1 < n_it) {
          ptx::mbar_wait(B.q_full + 8 * (s ^ 1), ((g + 1) >> 1) & 1);
          ptx::tc_fence_after();
          gemm_kk(tS, b16, C::kK, qn16, C::kQ);                          // (1) S^T(it+1)
          ptx::mma_commit(B.s_full);
        }
        FA_TRACE_EV(tr, 5);
        ptx::mbar_wait(B.ds_ready, g & 1);
        ptx::tc_fence_after();
        FA_TRACE_EV(tr, 6);
        // dQ first: it heads the longest dependency chain of the loop (dQ -> drain -> dP^T(it+1) -> dS
        // -> dQ(it+1), through the shared tensor-memory columns); dK only feeds its own accumulator
        if constexpr (kMode == 0) {
#pragma unroll
          for (int ks = 0; ks < kT / 16; ++ks)                             // (5) dQ = dS K
            ptx::mma_ss(tDP, ptx::smem_desc_at(hi_mn, b16, C::kDS + ks * 2048),
                        ptx::smem_desc_at(hi_mn, b16, C::kK + ks * 2048), id_mn_mn, ks > 0);
          ptx::mma_commit(B.dq_full);
        }
#pragma unroll
        for (int ks = 0; ks < kT / 16; ++ks) {                             // (4) dK += dS^T Q
          const uint32_t aoff = (ks / 4) * C::kSub + (ks % 4) * 32;
          ptx::mma_ss(tDK, ptx::smem_desc_at(hi_k, b16, C::kDS + aoff), ptx::smem_desc_at(hi_mn, q16, C::kQ + ks * 2048),
                      id_k_mn, it > 0 || ks > 0);
        }
        ptx::mma_commit(B.q_empty + 8 * s);
      }
      ptx::mma_commit(B.acc_done);
      g0 += n_it;
      ++nv;
      }   // items
    }
  } else if (warp >= kDrainWarp0) {
    // =========================== dQ drain warpgroup ======================================
    // register split (16 warps x 128 = 2048 per lane): without dropout the drain warps hold a whole
    // dQ tile (208), with dropout the compute warps need the room for the keep-bit exchange
    if constexpr (kDrainRegs < 128) ptx::setmaxnreg_dec<kDrainRegs>();
    else ptx::setmaxnreg_inc<kDrainRegs>();
    const int dw = warp - kDrainWarp0;                 // == warp % 4: TMEM lane quadrant
    const int r = dw * 32 + lane;                      // query row inside the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)(dw * 32) << 16;
    const bool leader = (warp == kDrainWarp0 && lane == 0);
    uint32_t cc = 0;                                   // running 16 KB block counter (staging ring of 2)
    const CUtensorMap *mdk = &tmap_dk, *mdv = &tmap_dv;
    FA_TRACE_DECL(trd, 1, tr_cta && leader);
    int g0 = 0, nv = 0, mapped_b = -1, set = 0;
    for (int n = 0;; ++n) {
    const int L = next_item(n);
    if (L < 0) break;
    const Item I = decode(L);
    if (!I.valid) continue;
    FA_ITEM_LOCALS(I);
    if (a.cu_q && warp == kDrainWarp0 && bidx != mapped_b) {
      // packed inputs: rows past this sequence's end belong to the next sequence -- the dK / dV stores are clipped by
      // maps bounded to this one.  The stores of the previous sequence must be over before its set is rewritten.
      if (lane == 0) ptx::tma_store_wait<0>();
      __syncwarp();
      const uint32_t sm = ptx::smid();
      if (sm >= (uint32_t)ptx::kMaxSm) __trap();
      const uint32_t d0 = sbase + C::kDesc;
      uint8_t *gen0 = smem_raw + C::kDesc;
      const uint32_t out_rows = kMirror ? (uint32_t)(qrow0 + sq) : (uint32_t)(krow0 + sk);   // mirrored: the map is dq's
      mdk = ptx::patch_rows(tmap_dk, &g_bwd_slots[sm][set][4], d0 + 512, gen0 + 512, out_rows, lane);
      mdv = ptx::patch_rows(tmap_dv, &g_bwd_slots[sm][set][5], d0 + 640, gen0 + 640, out_rows, lane);
      mapped_b = bidx;
      set ^= 1;
    }
    if constexpr (kMirror) mirror_item(std::integral_constant<int, 32>{}, 96, I, g0);   // my share of the score work
    for (int it = 0; kMode == 0 && it < n_it; ++it) {
      const int hq = hk * group + it / n_i, i = q_tile_of(it);
      // dq_accum layout: [b][h][query tile][D/32][128 rows][32 floats], 16-byte chunks XOR-swizzled by row
      float *gdst = a.dq_accum + ((((size_t)bidx * a.h + hq) * a.nqt + i) * a.nblk) * 4096;
      FA_TRACE_EV(trd, 20);
      ptx::mbar_wait(B.dq_full, (g0 + it) & 1);
      ptx::tc_fence_after();
      FA_TRACE_EV(trd, 21);
      // kEarlyDrain: the whole dQ tile -> registers first, so dP^T(it+1) may reuse the TMEM columns at once
      uint32_t v[kEarlyDrain ? D / 32 : 1][32];
      if constexpr (kEarlyDrain) {
#pragma unroll
        for (int c = 0; c < D / 32; ++c) ptx::tmem_ld_x32(tDP + lane_sel + c * 32, v[c]);
        ptx::tmem_wait_ld();
        ptx::tc_fence_before();
        ptx::mbar_arrive(B.dq_drained);
      }
      FA_TRACE_EV(trd, 22);
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        if (c >= a.nblk) break;                        // head_size < D: the last blocks are all zero
        uint32_t(&vc)[32] = v[kEarlyDrain ? c : 0];
        if constexpr (!kEarlyDrain) {
          ptx::tmem_ld_x32(tDP + lane_sel + c * 32, vc);
          ptx::tmem_wait_ld();
          if (c == D / 32 - 1) {                       // TMEM tile fully read: dP^T(it+1) may overwrite it
            ptx::tc_fence_before();
            ptx::mbar_arrive(B.dq_drained);
          }
        }
        const uint32_t buf = sbase + C::kDQ + (cc & 1) * 16384;
        if (leader && cc >= 2) ptx::tma_store_wait_read<1>();   // the block that used `buf` has been read
        ptx::named_bar_sync(1, 128);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          st_shared_v4(buf + r * 128 + ((k ^ (r & 7)) << 4), vc[4 * k], vc[4 * k + 1], vc[4 * k + 2], vc[4 * k + 3]);
        ptx::fence_proxy_async_smem();
        ptx::named_bar_sync(1, 128);
        if (leader) {
          bulk_reduce_add_f32(gdst + c * 4096, buf, 16384);
          ptx::tma_store_commit();
        }
        ++cc;
      }
    }
    // ---- epilogue: dK_j, dV_j -> global ----------------------------------------------------
    // A drain thread owns one ROW of dK / dV (TMEM lane = key row).  The tiles leave like the dQ blocks do:
    // 64 columns x 128 rows are parked in a 16 KB staging block in the tensor map's 128-byte-swizzled box
    // layout and ONE bulk tensor store sends the box (rows past the sequence end and columns past head_size
    // are clipped by the map) -- while the compute warps are already in the next item.  (They used to stage
    // both tiles in the Q ring and write them out themselves: ~5000 cycles per item on a store port the dQ
    // tiles keep 64 % busy.  One small bulk copy per row was measured too: ~18 cycles per copy, 18800 per item.)
    ptx::mbar_wait(B.acc_done, nv & 1);
    ptx::tc_fence_after();
    FA_TRACE_EV(trd, 23);
#pragma unroll
    for (int which = kMirror ? 1 : 0; which < 2; ++which) {        // mirrored: only the dK columns are used (they hold dQ)
      const float mul = which ? a.scale * a.rp_dropout : a.rp_dropout;
#pragma unroll
      for (int c = 0; c < D / 64; ++c) {
        const uint32_t buf = sbase + C::kDQ + (cc & 1) * 16384;
        if (leader && cc >= 2) ptx::tma_store_wait_read<1>();   // the block that used `buf` has been read
        ptx::named_bar_sync(1, 128);
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          uint32_t v[32];
          if (n_it > 0) {
            ptx::tmem_ld_x32((which ? tDK : tDV) + lane_sel + c * 64 + hf * 32, v);
            ptx::tmem_wait_ld();
          } else {
#pragma unroll
            for (int x = 0; x < 32; ++x) v[x] = 0;
          }
          if (which == 1 && c == D / 64 - 1 && hf == 1) {     // dK / dV read: the next item's MMAs may overwrite them
            ptx::tc_fence_before();
            ptx::mbar_arrive(B.epi_tmem_free);
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
          if (c * 64 < a.d) ptx::tma_store_4d(which ? mdk : mdv, buf, c * 64, hk, (kMirror ? qrow0 : krow0) + j * kT, bcoord);
          ptx::tma_store_commit();
        }
        ++cc;
      }
    }
    FA_TRACE_EV(trd, 24);
    if constexpr (kBiasTma) {
      if (a.mask_tma && leader) {          // the staging blocks have been read by the stores: the next item's bias tile may land
        ptx::tma_store_wait_read<0>();
        ptx::mbar_arrive(B.epi_smem_free);
      }
    }
    g0 += n_it;
    ++nv;
    }   // items
    if (leader) ptx::tma_store_wait<0>();
  } else {
    // =========================== compute warps ============================================
    ptx::setmaxnreg_inc<kComputeRegs>();
    const int w = warp >> 2;                           // which 64 query columns
    const int n = (warp & 3) * 32 + lane;              // key row inside the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    const float c2 = a.scale_log2;
    uint32_t add7 = 0;
    bool t_msb = false;
    if constexpr (kDropout) {
      add7 = (0x80u - (a.drop_threshold & 0x7fu)) * 0x01010101u;
      t_msb = a.drop_threshold >= 128u;
    }

    // Dropout keep bits for my 64 (query, key) pairs per iteration (csrc/philox_spec.h): the lanes of a
    // half-warp share the key block, so each lane draws the 16 bytes of ONE query row per 16-row block
    // (four calls) and the rows are exchanged.  The draws do not depend on S^T: iteration it + 1's four
    // calls are stepped under iteration it's exponentials (five rounds per 8 MUFU, as in the forward).
    FA_TRACE_DECL(trc, 2 + w, tr_cta && (warp & 3) == 0 && lane == 0);
    int g0 = 0;
    for (int item_n = 0;; ++item_n) {
    const int L = next_item(item_n);
    if (L < 0) break;
    const Item I = decode(L);
    if (!I.valid) continue;
    if constexpr (kMirror) {                           // the mirrored pass: columns [0, 48) and [48, 96) (the drain warps take the rest)
      mirror_item(std::integral_constant<int, 48>{}, w * 48, I, g0);
      g0 += I.n_it;
      continue;
    }
    FA_ITEM_LOCALS(I);
    const int n_abs = j * kT + n;
    const bool kv_valid = n_abs < sk;
    const bool ragged_k = j * kT + kT > sk;            // this item's key tile runs past the sequence end
    uint32_t pc[4][4];
    auto ctr_init = [&](int it_) {
      const int hq_ = hk * group + it_ / n_i, i_ = q_tile_of(it_);
      const uint32_t bh_ = (uint32_t)((a.batch_base + bidx) * a.h + hq_);
#pragma unroll
      for (int blk = 0; blk < 4; ++blk) {
        pc[blk][0] = (uint32_t)(i_ * kT + w * 64 + blk * 16 + (lane & 15));
        pc[blk][1] = (uint32_t)(n_abs >> 4); pc[blk][2] = bh_; pc[blk][3] = a.ctr3;
      }
    };
    if constexpr (kDropout) {
      if (n_it > 0) {
        ctr_init(0);
#pragma unroll
        for (int r = 0; r < FA_PHILOX_ROUNDS; ++r)
#pragma unroll
          for (int blk = 0; blk < 4; ++blk) fa_philox_round(pc[blk], a.key0, a.key1, r);
      }
    }

    for (int it = 0; it < n_it; ++it) {
      const int hq = hk * group + it / n_i, i = q_tile_of(it);
      const int g = g0 + it, s = g & 1;
      FA_TRACE_EV(trc, 10);
      const uint32_t stat = sbase + C::kStat + s * 1024 + w * 256;   // lse2 / D of my 64 query columns
      const int q_base = i * kT + w * 64;

      // dropout keep bits for my 64 (query, key) pairs: lanes of a half-warp share the key
      // block, so each lane draws the 16 bytes of ONE query row and the rows are exchanged.
      uint32_t keep_lo = 0xffffffffu, keep_hi = 0xffffffffu;
      if constexpr (kDropout) {
        // pc[blk] holds the finished Philox words of query row q_base + 16 blk + lane % 16 for THIS iteration:
        // drawn whole before the loop (it == 0) or stepped under the previous iteration's exponentials
        uint32_t w01 = 0, w23 = 0;        // bit 16*b + k: keep(query q_base + 16*b' + lane%16, key k of my block)
#pragma unroll
        for (int blk = 0; blk < 4; ++blk) {
          fa_philox4_t rnd;
#pragma unroll
          for (int x = 0; x < 4; ++x) rnd.v[x] = pc[blk][x];
          const uint32_t mine = keep_bits16(rnd, add7, t_msb);
          if (blk < 2) w01 |= mine << (16 * blk);
          else w23 |= mine << (16 * (blk - 2));
        }
        // 16x16 bit transposes inside each half-warp (four at once, two per word): butterfly over
        // lane bits 8, 4, 2, 1 -- 8 shuffles instead of the 64 a gather per (query, key) pair costs
        bit_transpose_stage<8>(w01, w23, lane);
        bit_transpose_stage<4>(w01, w23, lane);
        bit_transpose_stage<2>(w01, w23, lane);
        bit_transpose_stage<1>(w01, w23, lane);
        keep_lo = w01;                    // bit t: keep(query q_base + t, my key)
        keep_hi = w23;
      }

      // attn_mask (reference flash_attn.h:98-99): the 64 bias values of my key and the iteration's query rows do not
      // depend on S^T either -- fetched before the wait for it, so the loads' latency is spent waiting anyway
      // (lanes hold consecutive keys: each load instruction reads 64 contiguous bytes of one mask row)
      uint16_t mbias[kMask ? 64 : 1];
      if (kBiasTma && a.mask_tma) {
        ptx::mbar_wait(B.bias_full, g & 1);
        const uint16_t *bt = reinterpret_cast<const uint16_t *>(smem_raw + C::kDQ) + (size_t)(w * 64) * kT + n;   // [query][key]
        const bool rows_ragged = q_base + 64 > sq;         // warp-uniform
#pragma unroll
        for (int c = 0; c < 64; ++c) {
          const uint16_t vb = bt[c * kT];
          mbias[c] = (kv_valid && (!rows_ragged || q_base + c < sq)) ? vb : (uint16_t)0;   // (padding may hold anything)
        }
        ptx::mbar_arrive(B.bias_empty);
      } else if constexpr (kMask) {
        const uint16_t *mp = static_cast<const uint16_t *>(a.mask) +
            (((size_t)(a.mask_b == 1 ? 0 : bidx) * a.mask_h + (a.mask_h == 1 ? 0 : hq)) * a.mask_sq + q_base) * (size_t)a.mask_sk + n_abs;
#pragma unroll
        for (int c = 0; c < 64; ++c)
          mbias[c] = (kv_valid && q_base + c < sq) ? __ldg(mp + (size_t)c * a.mask_sk) : (uint16_t)0;
      }
      ptx::mbar_wait(B.q_full + 8 * s, (g >> 1) & 1);    // LSE / D staged
      FA_TRACE_EV(trc, 11);
      ptx::mbar_wait(B.s_full, g & 1);
      ptx::tc_fence_after();
      FA_TRACE_EV(trc, 12);
      float p[64];
      {
        uint32_t(&r0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&p[0]);
        uint32_t(&r1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&p[32]);
        ptx::tmem_ld_x32(tS + lane_sel + w * 64, r0);
        ptx::tmem_ld_x32(tS + lane_sel + w * 64 + 32, r1);
        ptx::tmem_wait_ld();
      }
      FA_TRACE_EV(trc, 30);
      const bool diag = a.causal && (j * kT + kT - 1 > q_base + shift);   // warp-uniform
      const int lim = n_abs - shift;                     // causal: visible iff q >= lim
      float cs = c2;
      if constexpr (kMask) {                                      // s <- s * scale_log2 + bias * log2 e
#pragma unroll
        for (int c = 0; c < 64; ++c) {
          const uint16_t mb = mbias[c];
          const float bias = kBF16 ? __uint_as_float((uint32_t)mb << 16) : __half2float(*reinterpret_cast<const __half *>(&mb));
          p[c] = fmaf(p[c], c2, bias * kLog2e);
        }
        cs = 1.f;
      }
      const float2 c2v = make_float2(cs, cs);
      if constexpr (kDropout) ctr_init(it + 1);           // (one wasted draw after the last iteration)
#pragma unroll
      for (int c = 0; c < 64; c += 4) {
        const float4 nl = ld_shared_f32x4(stat + 4 * c);                  // -lse2 of 4 query columns
        float2 x = __ffma2_rn(make_float2(p[c], p[c + 1]), c2v, make_float2(nl.x, nl.y));
        float2 y = __ffma2_rn(make_float2(p[c + 2], p[c + 3]), c2v, make_float2(nl.z, nl.w));
        p[c] = ptx::ex2(x.x); p[c + 1] = ptx::ex2(x.y);
        if (kBwdEmuEvery && (c / 4) % kBwdEmuEvery == 0) {   // this pair on the FMA pipe (ptx::ex2_poly2)
          y = ptx::ex2_poly2(y);
          p[c + 2] = y.x; p[c + 3] = y.y;
        } else {
          p[c + 2] = ptx::ex2(y.x); p[c + 3] = ptx::ex2(y.y);
        }
        if constexpr (kDropout) {
          if (c % 8 == 4) {               // a group = 8 exponentials + 5 Philox rounds, round-robin over the calls
            const int g = c / 8;
#pragma unroll
            for (int k = kPhiloxSteps * g / 8; k < kPhiloxSteps * (g + 1) / 8; ++k)
              fa_philox_round(pc[k % 4], a.key0, a.key1, k / 4);
            // nvvm would sink the rounds to their use and ptxas below the last exponential (a warp issues in
            // order: they would run after the MUFU stream, not under it): laundering pins them to this
            // group, __syncwarp() ends ptxas's scheduling block
#pragma unroll
            for (int blk = 0; blk < 4; ++blk)
#pragma unroll
              for (int x2 = 0; x2 < 4; ++x2) pc[blk][x2] = ptx::opaque(pc[blk][x2]);
            if (kBwdSyncGroups && g % kBwdSyncGroups == kBwdSyncGroups - 1) __syncwarp();
          }
        }
      }
      FA_TRACE_EV(trc, 31);
      if (diag || ragged_k) {                            // CTA/warp-uniform: only diagonal and last key tiles
#pragma unroll
        for (int c = 0; c < 64; ++c)
          if (!kv_valid || (diag && q_base + c < lim)) p[c] = 0.f;
      }
      {
        uint32_t pk[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) pk[c] = pack2(p[2 * c], p[2 * c + 1], kBF16);
        if constexpr (kDropout) {
#pragma unroll
          for (int hw = 0; hw < 2; ++hw) {
            uint32_t wk[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) wk[k] = (hw ? keep_hi : keep_lo) << (7 - k);
#pragma unroll
            for (int c = 0; c < 16; ++c) pk[16 * hw + c] &= pair_mask(wk, c);
          }
        }
        FA_TRACE_EV(trc, 32);
        ptx::tmem_st_x32(tS + lane_sel + w * 64, pk);     // P^T half w lives in columns [64w, 64w+32)
        ptx::tmem_wait_st();
        FA_TRACE_EV(trc, 33);
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(B.p_ready);
      FA_TRACE_EV(trc, 13);

      ptx::mbar_wait(B.dp_full, g & 1);
      ptx::tc_fence_after();
      FA_TRACE_EV(trc, 14);
      // GEMMs (5),(4) of it-1 have read dS^T: q_empty is committed after the later of the two (dK).  Long
      // done by now, so waiting here (not after the math) costs nothing and lets each 16-column chunk of
      // dS^T leave for shared memory as soon as it is computed.
      if (g > 0) ptx::mbar_wait(B.q_empty + 8 * ((g - 1) & 1), ((g - 1) >> 1) & 1);
      FA_TRACE_EV(trc, 16);
      {
        // dP^T comes in four 16-column chunks; chunk q + 1 is in flight while chunk q is worked on
        const uint32_t rowp = sbase + C::kDS + w * C::kSub + n * 128;
        uint32_t dpb[2][16];
        ptx::tmem_ld_x16(tDP + lane_sel + w * 64, dpb[0]);
        ptx::tmem_wait_ld();
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          uint32_t(&dp)[16] = dpb[q4 & 1];
          if (q4 < 3) ptx::tmem_ld_x16(tDP + lane_sel + w * 64 + (q4 + 1) * 16, dpb[(q4 + 1) & 1]);
          uint32_t ds[8];
#pragma unroll
          for (int c = 0; c < 16; c += 4) {
            const int cq = q4 * 16 + c;
            float2 g01 = ptx::u2f2(dp[c], dp[c + 1]), g23 = ptx::u2f2(dp[c + 2], dp[c + 3]);
            if constexpr (kDropout) {
              const uint32_t kb = cq < 32 ? keep_lo >> cq : keep_hi >> (cq - 32);
              g01.x = (kb & 1u) ? g01.x : 0.f;
              g01.y = (kb & 2u) ? g01.y : 0.f;
              g23.x = (kb & 4u) ? g23.x : 0.f;
              g23.y = (kb & 8u) ? g23.y : 0.f;
            }
            const float4 nd = ld_shared_f32x4(stat + 512 + 4 * cq);        // -D of 4 query columns
            const float2 e01 = __fmul2_rn(make_float2(p[cq], p[cq + 1]), __fadd2_rn(g01, make_float2(nd.x, nd.y)));
            const float2 e23 = __fmul2_rn(make_float2(p[cq + 2], p[cq + 3]), __fadd2_rn(g23, make_float2(nd.z, nd.w)));
            ds[c / 2] = pack2(e01.x, e01.y, kBF16);
            ds[c / 2 + 1] = pack2(e23.x, e23.y, kBF16);
          }
#pragma unroll
          for (int k = 0; k < 2; ++k)
            st_shared_v4(rowp + (((2 * q4 + k) ^ (n & 7)) << 4), ds[4 * k], ds[4 * k + 1], ds[4 * k + 2], ds[4 * k + 3]);
          if (q4 < 3) ptx::tmem_wait_ld();
        }
      }
      FA_TRACE_EV(trc, 15);
      ptx::fence_proxy_async_smem();
      ptx::tc_fence_before();
      ptx::mbar_arrive(B.ds_ready);
      FA_TRACE_EV(trc, 17);
    }

    // (dK_j / dV_j are written out by the drain warps, see above: the compute warps go straight to the next item)
    FA_TRACE_EV(trc, 18);
    g0 += n_it;
    }   // items
  }

  // ---- teardown -------------------------------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}

}  // namespace

namespace {

// dq = softmax_scale * dq_accum, un-swizzling the [tile][D/32][128][32] scratch layout.
template <int D, bool kBF16>
__global__ void __launch_bounds__(256)
dq_convert_kernel(const float *__restrict__ acc, void *__restrict__ dq, const int32_t *__restrict__ cu_q,
                  int sq, int h, int nqt, float scale, int d, int nblk) {
  const int i = blockIdx.x, hq = blockIdx.y, bidx = blockIdx.z;
  size_t tok0 = (size_t)bidx * sq;
  if (cu_q) {
    tok0 = (size_t)cu_q[bidx];
    sq = cu_q[bidx + 1] - cu_q[bidx];
    if (i * 128 >= sq) return;
  }
  const float *src = acc + ((((size_t)bidx * h + hq) * nqt + i) * nblk) * 4096;
  const int kVecPerRow = d / 8;                        // 16-byte output vectors per row
  for (int idx = threadIdx.x; idx < 128 * kVecPerRow; idx += 256) {
    const int r = idx / kVecPerRow, vcol = idx % kVecPerRow;
    const int row = i * 128 + r;
    if (row >= sq) continue;
    const int d0 = vcol * 8, c = d0 / 32, k = (d0 % 32) / 4;       // two 16-byte chunks k, k+1
    const float *rowp = src + (size_t)c * 4096 + r * 32;
    const float4 x = *reinterpret_cast<const float4 *>(rowp + ((k ^ (r & 7)) << 2));
    const float4 y = *reinterpret_cast<const float4 *>(rowp + (((k + 1) ^ (r & 7)) << 2));
    uint4 o;
    o.x = pack2(x.x * scale, x.y * scale, kBF16);
    o.y = pack2(x.z * scale, x.w * scale, kBF16);
    o.z = pack2(y.x * scale, y.y * scale, kBF16);
    o.w = pack2(y.z * scale, y.w * scale, kBF16);
    uint16_t *dst = static_cast<uint16_t *>(dq) + ((tok0 + row) * h + hq) * d + d0;
    *reinterpret_cast<uint4 *>(dst) = o;
  }
}

std::atomic<SchedSlot *> g_bwd_sched_addr[kMaxCachedDevices];   // device address of g_bwd_sched, per device
std::atomic<unsigned> g_bwd_ticket{0};

// bias rows of whole 16-byte units from a 16-byte aligned base: what a TMA tensor map needs
bool mask_tma_ok(const Params &p) {
  return p.mask && p.mask_sk % 8 == 0 && (reinterpret_cast<uintptr_t>(p.mask) & 15) == 0;
}

template <int D, bool kBF16, bool kDropout, bool kMask, int kMode>
const char *launch_tc_bwd_t(const Params &p, cudaStream_t stream) {
  using C = Cfg<D>;
  CUtensorMap tq, tk, tv, tdo;
  const char *e;
  // packed: one batch entry whose row count each CTA replaces (ptx::patch_rows)
  const uint64_t nb = p.cu_q ? 1 : (uint64_t)p.b;
  if ((e = make_tmap_rows(&tq, p.q, kBF16, p.d, p.h, p.sq, nb, kT))) return e;
  if ((e = make_tmap_rows(&tdo, p.dout, kBF16, p.d, p.h, p.sq, nb, kT))) return e;
  if ((e = make_tmap_rows(&tk, p.k, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
  if ((e = make_tmap_rows(&tv, p.v, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
  CUtensorMap tdk, tdv;      // the dK / dV tiles leave through bulk tensor stores (mode 2: dQ, through the dK map)
  if (kMode == 2) {
    if ((e = make_tmap_rows(&tdk, p.dq, kBF16, p.d, p.h, p.sq, nb, kT))) return e;
    tdv = tdk;
  } else {
    if ((e = make_tmap_rows(&tdk, p.dk, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
    if ((e = make_tmap_rows(&tdv, p.dv, kBF16, p.d, p.hk, p.sk, nb, kT))) return e;
  }

  BwdArgs a;
  a.dk = p.dk; a.dv = p.dv; a.dq_accum = p.dq_accum; a.lse = p.lse; a.softmax_d = p.softmax_d;
  a.cu_q = p.cu_q; a.cu_k = p.cu_k;
  a.b = p.b; a.sq = p.sq; a.sk = p.sk; a.sq_rounded = p.sq_rounded; a.h = p.h; a.hk = p.hk;
  a.nqt = (p.sq + kT - 1) / kT;
  a.batch_base = p.batch_base;
  a.d = p.d;
  a.nblk = (p.d + 31) / 32;
  a.mask = p.mask; a.mask_b = (int)p.mask_b; a.mask_h = (int)p.mask_h; a.mask_sq = (int)p.mask_sq; a.mask_sk = (int)p.mask_sk;
  a.scale = p.scale; a.scale_log2 = p.scale * kLog2e; a.rp_dropout = p.rp_dropout;
  a.keep_prob = kDropout ? 1.f - p.p_dropout : 1.f;
  a.drop_threshold = p.drop_threshold;
  a.key0 = (uint32_t)p.seed ^ (uint32_t)(p.offset >> 32);
  a.key1 = (uint32_t)(p.seed >> 32);
  a.ctr3 = (uint32_t)p.offset;
  a.causal = p.causal ? 1 : 0;
  a.hk_kv = p.hk;

  CUtensorMap tbias = tq;    // (a valid map either way; only read when mask_tma is set)
  a.mask_tma = 0;
  if (kMask && kMode == 1 && mask_tma_ok(p)) {
    if ((e = make_tmap_bias(&tbias, p.mask, kBF16, (uint64_t)p.mask_sk, (uint64_t)p.mask_sq, (uint64_t)p.mask_h, (uint64_t)p.mask_b, kT, kT)))
      return e;
    a.mask_tma = 1;
  }
  auto kern = tc_bwd_kernel<D, kBF16, kDropout, kMask, kMode>;
  int dev = 0, sm_count = 0;
  if ((e = current_device(&dev, &sm_count))) return e;
  static std::atomic<uint64_t> smem_opt_in{0};   // per instantiation, one bit per device
  if ((e = ensure_dynamic_smem(kern, (int)C::kBytes, dev, smem_opt_in))) return e;
  a.nkt = (p.sk + kT - 1) / kT;
  a.n_items = a.nkt * p.hk * p.b;
  if (kMode == 2) {          // mirrored: items are (query tile, QUERY head, batch); the kernel reads them from nkt / hk
    a.nkt = a.nqt;
    a.hk = p.h;
    a.n_items = a.nqt * p.h * p.b;
  }
  a.persistent = a.n_items > sm_count ? 1 : 0;   // no more items than SMs: no counter needed
  a.sched = nullptr;
  int grid = a.n_items;
  if (a.persistent) {
    grid = sm_count;
    if ((e = take_sched_slot(g_bwd_sched, dev, g_bwd_sched_addr, g_bwd_ticket, &a.sched))) return e;
  }

  // (mode 0: dq_accum was zeroed by the preprocess kernel: tc_bwd_scratch_floats)
  kern<<<grid, kThreads, C::kBytes, stream>>>(tq, tk, tv, tdo, tdk, tdv, tbias, a);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) return cudaGetErrorString(err);
  if (kMode != 0) return nullptr;          // the two-launch deterministic backward never touches dq_accum
  dim3 cgrid(a.nqt, p.h, p.b);
  dq_convert_kernel<D, kBF16><<<cgrid, 256, 0, stream>>>(p.dq_accum, p.dq, p.cu_q, p.sq, p.h, a.nqt, p.scale * p.rp_dropout, a.d,
                                                         a.nblk);
  err = cudaGetLastError();
  return err == cudaSuccess ? nullptr : cudaGetErrorString(err);
}

}  // namespace

int trace_read_bwd(unsigned long long *dst, int role, int max_events) {
#ifdef FA_TRACE
  return trace::read_events(dst, role, max_events);
#else
  (void)dst; (void)role; (void)max_events;
  return 0;
#endif
}

// Whether launch_tc_bwd will add dQ tiles into dq_accum (then the preprocess kernel must zero it): the two-launch
// backward (num_splits == 1, attn_mask with TMA-able rows; head sizes up to 128) never touches the scratch.
// The caller's scratch pair as include/flash_attn.h sizes it: dq_accum >= b * sq_rounded * h * d_rounded floats with
// sq_rounded a multiple of 128 and d_rounded >= head_size rounded up to 32.  A call whose scratch is smaller (or NULL)
// still runs on the tcgen05 kernels: as the two-launch pair, which does not use dq_accum at all.
bool scratch_fits(const Params &p) {
  return p.dq_accum && p.sq_rounded % kT == 0 && p.sq_rounded >= p.sq && p.d_rounded >= (p.d + 31) / 32 * 32;
}

bool tc_bwd_uses_scratch(const Params &p) {
  if (!scratch_fits(p)) return false;
  if (p.d > 128) return !(p.deterministic || p.mask);      // (wide heads: attn_mask always takes the two launches)
  return !(p.deterministic || (p.mask && mask_tma_ok(p)));
}

long long tc_bwd_scratch_floats(const Params &p) {
  return (long long)p.b * p.h * ((p.sq + kT - 1) / kT) * kT * ((p.d + 31) / 32) * 32;
}

bool tc_bwd_supported(const Params &p) {
  if (p.d <= 0 || p.d > 256 || p.d % 8) return false;   // d is padded to 64, 128 or 256 by TMA zero fill
  // (any finite softmax_scale, zero and negative included: P is recomputed from the stored LSE, no running maximum is
  // tracked here -- the forward, which tracks one on raw scores, takes it through a negated MMA operand or, for zero, a 1e-30 stand-in)
  if (p.sq <= 0 || p.sk <= 0) return false;
  return true;                                          // (a scratch pair below the contract: two launches, see scratch_fits)
}

namespace {
// num_splits == 1: dK / dV, then dQ, no atomics anywhere (see kMode at the kernel)
template <int D, bool kBF16, bool kDropout, bool kMask>
const char *launch_det(const Params &p, cudaStream_t stream) {
  const char *e = launch_tc_bwd_t<D, kBF16, kDropout, kMask, 1>(p, stream);
  return e ? e : launch_tc_bwd_t<D, kBF16, kDropout, kMask, 2>(p, stream);
}
}  // namespace

const char *launch_tc_bwd(const Params &p_in, cudaStream_t stream) {
  Params p = p_in;
  if (!scratch_fits(p)) p.deterministic = true;         // no usable dq_accum: the pair that needs none
  if (p.d > 128) return launch_tc_bwd_wide(p, stream);
  const bool drop = p.p_dropout > 0.f;
#define FA_MASKED(DD, BF, DR)                                                              \
  (p.mask ? ((p.deterministic || mask_tma_ok(p)) ? launch_det<DD, BF, DR, true>(p, stream)   \
                                                 : launch_tc_bwd_t<DD, BF, DR, true, 0>(p, stream)) \
          : p.deterministic ? launch_det<DD, BF, DR, false>(p, stream)                     \
                            : launch_tc_bwd_t<DD, BF, DR, false, 0>(p, stream))
#define FA_DISPATCH(DD)                                                                          \
  if (p.is_bf16) return drop ? FA_MASKED(DD, true, true)                          \
                             : FA_MASKED(DD, true, false);                        \
  return drop ? FA_MASKED(DD, false, true) : FA_MASKED(DD, false, false);
  if (p.d > 64) { FA_DISPATCH(128) }
  FA_DISPATCH(64)
#undef FA_DISPATCH
#undef FA_MASKED
}

}  // namespace fa
