// fa_tc_fwd.cu -- FlashAttention-2 forward for sm_100a: TMA -> shared memory,
// tcgen05.mma with accumulators in tensor memory, online softmax in registers.
// Implements flash_attn_fwd / flash_attn_varlen_fwd (reference flash_attn.h:10-66)
// for head_size 64/128, fp16/bf16, causal or not, GQA, Philox dropout.
//
// One CTA owns 256 query rows of one (batch, head): two 128-row tiles that
// ping-pong on the tensor pipe.  Twelve warps:
//   warps 0-3   softmax of tile 0 (thread r owns query row r: TMEM lane r)
//   warps 4-7   softmax of tile 1
//   warp  8     tcgen05.mma issue (one lane); allocates/frees tensor memory
//   warp  9     TMA producer (one lane)
//   warps 10-11 idle (they only complete the warpgroup for setmaxnreg)
// Tensor memory (512 columns x 128 lanes x 32 bit):
//   [0,128) S0 | [128,256) S1 | [256,256+D) O0 | [384,384+D) O1
//   P (bf16 pairs) overwrites columns [kPOff, kPOff+64) of its S tile.
// Per KV tile j and query tile t, in tensor-pipe issue order:
//   S_t = Q_t K_j^T            (A, B from shared memory, both K-major)
//   softmax warps: S_t -> regs, mask, running max / sum, P_t -> TMEM
//   O_t (+)= P_t V_j           (A from tensor memory, B MN-major from shared)
// The running max is updated lazily: O_t is rescaled (by the softmax warps
// themselves) only when the max grew by more than 2^8, which keeps P <= 256 in
// bf16 and takes the rescale off the steady-state critical path.
#include <type_traits>
#include "fa_common.cuh"
#include "ptx_sm100.cuh"
#include "tma_host.cuh"
#include "fa_trace.cuh"

namespace fa {

namespace {

constexpr int kTileM = 128;        // query rows per tile (= TMEM lanes)
constexpr int kTileN = 128;        // keys per KV tile
constexpr int kThreads = 384;
constexpr int kMmaWarp = 8, kLoadWarp = 9;
constexpr uint32_t kPOff = 32;     // P's first column inside its S tile
constexpr float kRescaleThreshold = 8.0f;  // in log2 units
#ifndef FA_EMU_PER8
#define FA_EMU_PER8 0
#endif
// Exponentials per 8 computed on the FMA pipe (ptx::ex2_poly2) instead of MUFU.  Measured on B200
// (S4096 D128, build.py --variant): 0 -> 1075, 2 -> 1057, 4 -> 1018, 6 -> 981 TFLOP/s causal: the
// softmax is bound by the S -> P -> PV -> QK dependency chain and by issue slots, not by MUFU.
constexpr int kEmuPer8 = FA_EMU_PER8;
#ifndef FA_FWD_SYNC_GROUPS
#define FA_FWD_SYNC_GROUPS 2
#endif
// Dropout: a scheduling block of the softmax loop = this many groups of (8 exponentials + 3-4 Philox rounds; 5 when
// the generator had 10 rounds, which is what the figures below were measured with).
// ptxas orders instructions freely inside a block and a warp issues in order, so the block size decides how
// well the integer work hides under the MUFU stream.  Measured on B200 (C3 forward, dropout 0.1):
// whole tile 710, 4 groups 778, 3 groups 811, 2 groups 818, 1 group 747 TFLOP/s (660 with the draws after S arrived).
constexpr int kSyncGroups = FA_FWD_SYNC_GROUPS;
// The draw-ahead loop spreads 4 calls x FA_PHILOX_ROUNDS rounds over the 8 groups of a 64-column half.
constexpr int kPhiloxSteps = 4 * FA_PHILOX_ROUNDS;
#ifndef FA_FWD_APPLY_IN_GROUP
#define FA_FWD_APPLY_IN_GROUP 0
#endif
// AND the keep masks inside the groups (1) or after the half's last exponential (0): 798 against 818 TFLOP/s
// (tests/cuda/ab_bench.py, interleaved rounds in one process).
constexpr bool kApplyInGroup = FA_FWD_APPLY_IN_GROUP != 0;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

template <int D>
struct Cfg {
  // D = 256 (head sizes 136..256): one 128-row query tile per CTA -- O alone takes 256 of the 512 tensor-
  // memory columns -- and single K / V buffers (Q + K + V = 192 KB), so the tensor pipe waits for the
  // softmax: roughly half the rate of the two-tile kernels, two orders above the all-shapes family.
  static constexpr int kStages = D == 128 ? 2 : D == 64 ? 4 : 1;   // K and V ring depth
  static constexpr int kQTiles = D == 256 ? 1 : 2;            // query tiles (softmax warpgroups at work) per CTA
  static constexpr uint32_t kTileBytes = kTileM * D * 2;      // one Q/K/V tile
  static constexpr uint32_t kSubBytes = kTileM * 128;         // one [128 x 64] swizzled sub-tile
  static constexpr uint32_t kSmemQ = 0;
  static constexpr uint32_t kSmemK = kQTiles * kTileBytes;
  static constexpr uint32_t kSmemV = kSmemK + kStages * kTileBytes;
  static constexpr uint32_t kSmemBar = kSmemV + kStages * kTileBytes;
  static constexpr uint32_t kNumBars = 1 + 4 * kStages + 8 + 7;
  static constexpr uint32_t kSmemDesc = kSmemBar + 512;       // 3 x 128 B tensor-map staging (packed inputs)
  static constexpr uint32_t kSmemO = kSmemDesc + 3 * 128;     // epilogue staging: 4 KB (32 rows x 64 columns) per softmax warp
  static constexpr uint32_t kSmemBytesNoStage = kSmemO + 1024;                // + alignment slack
  static constexpr uint32_t kSmemBytes = kSmemO + kQTiles * 4 * 4096 + 1024;
  static_assert(kSmemBytes <= 232448, "over the 227 KB a CTA may ask for");
  static constexpr uint32_t smem_bytes(bool stage_o) { return stage_o ? kSmemBytes : kSmemBytesNoStage; }
  static_assert(kNumBars * 8 + 16 <= 512, "barrier block (+ tmem slot, item queue) overflows");
  static_assert(kSmemBytesNoStage > 116 * 1024, "one CTA per SM is what makes the per-SM tensor-map slots private");
};

// O leaves through a per-warp shared-memory stage as whole 128-byte lines: epilogue 5100 -> 2700 cycles per
// tile, S2048 causal 797 -> 850 TFLOP/s.  Not in the dropout builds (measured slower there).
constexpr bool stage_o(bool dropout) {
#ifdef FA_FWD_STAGE_O_ALWAYS
  (void)dropout;
  return true;
#else
  return !dropout;
#endif
}

__device__ SchedSlot g_fwd_sched[kSchedSlots];       // item counters of the persistent launches (fa_common.cuh)

// Per-SM homes of the tensor maps a CTA bounds to its own sequence (packed inputs only).
__device__ ptx::TensorMapSlot g_fwd_slots[ptx::kMaxSm][2][3];   // two sets: the next sequence's maps are built while the previous ones are in use

struct Bars {
  uint32_t q_full, k_full, k_empty, v_full, v_empty, s_full, p_lo, p_hi, pv_done;  // first of each group
  uint32_t q_empty, o_free /*[2]*/, sched_full /*[2]*/, sched_empty /*[2]*/;
};

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}

__device__ __forceinline__ uint32_t pack_pair(float lo, float hi, bool bf16) {
  if (bf16) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
  }
  const __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t *>(&v);
}

// Turns 4 random bytes into two AND-masks for packed 16-bit pairs: byte k of
// `w` keeps element k iff byte < T (csrc/philox_spec.h).  SWAR compare on the
// low 7 bits, the top bit folded in, then PRMT's sign-replicate mode expands
// the per-byte verdict (bit 7) to 16 bits.
struct DropConst {
  uint32_t add7;   // (0x80 - (T & 0x7f)) replicated
  bool t_msb;      // T >= 128
};
// bit 7 of each byte of the result: that byte of `w` is < T (the entry is kept)
__device__ __forceinline__ uint32_t keep_word(uint32_t w, const DropConst &dc) {
  const uint32_t ge7 = (w & 0x7f7f7f7fu) + dc.add7;           // bit7 of each byte: low7 >= T7
  return dc.t_msb ? ~(w & ge7) : ~(w | ge7);
}
__device__ __forceinline__ void expand_keep(uint32_t keep, uint32_t &m01, uint32_t &m23) {
  // prmt's selector bit 3 replicates the byte's sign; __byte_perm() masks it off, hence asm
  asm("prmt.b32 %0, %1, %1, 0x9988;" : "=r"(m01) : "r"(keep));
  asm("prmt.b32 %0, %1, %1, 0xbbaa;" : "=r"(m23) : "r"(keep));
}

}  // namespace

namespace {

struct TcFwdArgs {
  void *out;
  float *lse;
  const int32_t *cu_q, *cu_k;
  int b, sq, sk, h, hk, batch_base;
  const void *mask;      // additive bias [mask_b, mask_h, mask_sq, mask_sk] in the input dtype, or null; packed inputs:
  int mask_b, mask_h;    // mask_sq / mask_sk = max_seqlen_q / max_seqlen_k, entry (row, col) counted from the sequence's start
  int mask_sq, mask_sk;
  int d;                 // real head_size (<= D, multiple of 8); columns [d, D) are TMA zero fill
  float scale_log2;      // |softmax_scale| * log2(e)
  uint32_t qk_negate;    // softmax_scale < 0: the instruction descriptor's negate-A bit, so S = -(Q K^T) and the scale stays positive
  float rp_dropout;
  uint32_t drop_threshold;
  uint32_t key0, key1, ctr3;   // Philox key / counter word 3 (csrc/philox_spec.h)
  int causal;
  int nqb, n_items;      // query blocks per (batch, head); work items = nqb * h * b
  int persistent;        // 1: CTAs pull items from *sched until none are left; 0: item = blockIdx.x
  SchedSlot *sched;      // global item counter of this launch (self-resetting pool slot, fa_common.cuh)
  // O addressing of the staged epilogue, in elements: out + batch * o_batch + row * o_row + head * o_head
  // (dense default: sq*h*d, h*d, d; the GQA-packed decode view swaps the roles of heads and rows)
  long long o_batch, o_row, o_head;
  int use_emu;           // D = 64 builds: FMA-pipe exp2 on (long key sequences) or off (short ones: -8 % at S1024 causal)
};

// One unit of work: 256 (128 when D = 256) query rows of one (batch, head).
struct Item {
  int bidx, head, hk, q0, k0, sq, sk, bcoord, m0, shift, nkv, nmax;
  int ntile[2];
  bool has_tile1, valid;
};

template <int D, bool kBF16, bool kDropout, bool kMask>
__global__ void __launch_bounds__(kThreads, 1)
tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
              const __grid_constant__ CUtensorMap tmap_v, const TcFwdArgs a) {
  using C = Cfg<D>;
  constexpr bool kStageO = stage_o(kDropout);
  constexpr bool kMaskStage = kMask && kStageO;       // the bias tile borrows the epilogue's staging blocks
  extern __shared__ uint8_t smem_raw[];
  const uint32_t sbase = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t *sgen = smem_raw + (sbase - ptx::smem_u32(smem_raw));
  // warp index through a shuffle: the compiler then knows the role branches are warp-uniform, and with
  // elect.sync inside them it issues TMA / tcgen05.mma without a per-instruction election loop
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

  // ---- work items ------------------------------------------------------------
  // Linear order: query blocks fastest inside a (batch, head), heaviest (last) causal block first, so
  // the CTAs running at any moment share a handful of heads' K/V in L2.  With more items than SMs one
  // CTA per SM takes items from a global counter (a.sched) until they run out, and its TMA and MMA warps
  // run ahead into the next item while the softmax warps finish the current one.  Packed inputs: the
  // load warp re-bounds its tensor maps whenever the next item belongs to another sequence (patch_rows).
  auto decode = [&](int L) {
    Item it;
    const int hb = L / a.nqb, r = L % a.nqb;
    it.head = hb % a.h; it.bidx = hb / a.h;
    const int qblk = a.causal ? a.nqb - 1 - r : r;
    it.q0 = 0; it.k0 = 0; it.sq = a.sq; it.sk = a.sk; it.bcoord = it.bidx;
    if (a.cu_q) {
      it.q0 = a.cu_q[it.bidx]; it.k0 = a.cu_k[it.bidx];
      it.sq = a.cu_q[it.bidx + 1] - it.q0; it.sk = a.cu_k[it.bidx + 1] - it.k0;
      it.bcoord = 0;
    }
    it.m0 = qblk * C::kQTiles * kTileM;
    it.valid = it.m0 < it.sq;
    it.hk = it.head / (a.h / a.hk);
    it.shift = it.sk - it.sq;                 // causal: key j visible to row i iff j <= i + shift
    it.nkv = (it.sk + kTileN - 1) / kTileN;
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const int r0 = it.m0 + t * kTileM;
      int n = (r0 < it.sq && t < C::kQTiles) ? it.nkv : 0;
      if (a.causal && n) {
        const int last = min(r0 + kTileM, it.sq) - 1 + it.shift;   // last visible key of the tile
        n = last < 0 ? 0 : min(it.nkv, last / kTileN + 1);
      }
      it.ntile[t] = n;
    }
    it.nmax = max(it.ntile[0], it.ntile[1]);
    it.has_tile1 = C::kQTiles == 2 && it.m0 + kTileM < it.sq;
    return it;
  };

  // ---- barriers, tensor memory ----------------------------------------------
  const uint32_t bar0 = sbase + C::kSmemBar;
  Bars B;
  B.q_full = bar0;
  B.k_full = bar0 + 8;
  B.k_empty = B.k_full + 8 * C::kStages;
  B.v_full = B.k_empty + 8 * C::kStages;
  B.v_empty = B.v_full + 8 * C::kStages;
  B.s_full = B.v_empty + 8 * C::kStages;
  B.p_lo = B.s_full + 16;
  B.p_hi = B.p_lo + 16;
  B.pv_done = B.p_hi + 16;
  B.q_empty = B.pv_done + 16;
  B.o_free = B.q_empty + 8;
  B.sched_full = B.o_free + 16;
  B.sched_empty = B.sched_full + 16;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(sgen + C::kSmemBar + C::kNumBars * 8);
  volatile int *sched_item = reinterpret_cast<volatile int *>(tmem_slot + 2);   // [2]
  constexpr int kSoftmaxThreads = C::kQTiles * kTileM;

  if (threadIdx.x == 0) {
    ptx::mbar_init(B.q_full, 1);
    ptx::mbar_init(B.q_empty, 1);
    for (int s = 0; s < C::kStages; ++s) {
      ptx::mbar_init(B.k_full + 8 * s, 1);
      ptx::mbar_init(B.k_empty + 8 * s, 1);
      ptx::mbar_init(B.v_full + 8 * s, 1);
      ptx::mbar_init(B.v_empty + 8 * s, 1);
    }
    for (int t = 0; t < 2; ++t) {
      ptx::mbar_init(B.s_full + 8 * t, 1);
      ptx::mbar_init(B.p_lo + 8 * t, kTileM);
      ptx::mbar_init(B.p_hi + 8 * t, kTileM);
      ptx::mbar_init(B.pv_done + 8 * t, 1);
      ptx::mbar_init(B.o_free + 8 * t, kTileM);
      ptx::mbar_init(B.sched_full + 8 * t, 1);
      ptx::mbar_init(B.sched_empty + 8 * t, 1 + kSoftmaxThreads);   // MMA lane + every softmax thread
    }
    ptx::fence_mbar_init();
  }
  if (warp == kLoadWarp && lane == 0) {
    ptx::prefetch_tensormap(&tmap_q);
    ptx::prefetch_tensormap(&tmap_k);
    ptx::prefetch_tensormap(&tmap_v);
  }
  if (warp == kMmaWarp) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
#ifdef FA_TRACE
  const bool tr_cta = blockIdx.x == gridDim.x / 2;
#endif
  // Consumers of the item queue (MMA lane, softmax threads): item number n of this CTA, or -1 at the end.
  auto next_item = [&](int n) {
    const int slot = n & 1;
    ptx::mbar_wait(B.sched_full + 8 * slot, (n >> 1) & 1);
    const int L = sched_item[slot];
    ptx::mbar_arrive(B.sched_empty + 8 * slot);
    return L;
  };

  if (warp >= 8) {
    // ======================= producer warpgroup ================================
    ptx::setmaxnreg_dec<40>();
    if (warp == kLoadWarp) {
      // ---- scheduler + TMA producer (lane 0 issues; the whole warp builds tensor maps) ----------
      const CUtensorMap *mq = &tmap_q, *mk = &tmap_k, *mv = &tmap_v;
      int mapped_b = -1, set = 0;   // packed inputs: the sequence the current maps are bounded to
      int nq = 0;        // items that loaded Q so far (q_full / q_empty phases)
      int g = 0;         // KV tiles loaded so far (ring stage and phase)
      for (int n = 0;; ++n) {
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
        const Item it = decode(L);
        if (!it.valid || it.nmax == 0) continue;
        if (a.cu_q && it.bidx != mapped_b) {
          // packed inputs: rows past this sequence's end must read as zero, not as the next sequence.  Loads in
          // flight may still be reading the maps of the previous sequence: two sets of per-SM slots alternate, and
          // (a change of sequence is rare) every load issued so far is seen complete before a set is rewritten.
          if (lane == 0 && mapped_b >= 0) {
            if (nq >= 1) ptx::mbar_wait(B.q_full, (nq - 1) & 1);
            for (int t = max(0, g - C::kStages); t < g; ++t) {
              ptx::mbar_wait(B.k_full + 8 * (t % C::kStages), (t / C::kStages) & 1);
              ptx::mbar_wait(B.v_full + 8 * (t % C::kStages), (t / C::kStages) & 1);
            }
          }
          __syncwarp();
          const uint32_t sm = ptx::smid();
          if (sm >= (uint32_t)ptx::kMaxSm) __trap();
          const uint32_t d0 = sbase + C::kSmemDesc;
          uint8_t *g0 = sgen + C::kSmemDesc;
          mq = ptx::patch_rows(tmap_q, &g_fwd_slots[sm][set][0], d0, g0, (uint32_t)(it.q0 + it.sq), lane);
          mk = ptx::patch_rows(tmap_k, &g_fwd_slots[sm][set][1], d0 + 128, g0 + 128, (uint32_t)(it.k0 + it.sk), lane);
          mv = ptx::patch_rows(tmap_v, &g_fwd_slots[sm][set][2], d0 + 256, g0 + 256, (uint32_t)(it.k0 + it.sk), lane);
          mapped_b = it.bidx;
          set ^= 1;
        }
        if (lane == 0) {
          const int qrow = it.q0 * (a.cu_q ? 1 : 0) + it.m0;   // packed: absolute token row; dense: row in sequence
          const int krow = it.k0 * (a.cu_q ? 1 : 0);
          if (nq >= 1) ptx::mbar_wait(B.q_empty, (nq - 1) & 1);  // the previous item's S = Q K^T are all done
          ptx::mbar_arrive_expect_tx(B.q_full, (it.has_tile1 ? 2 : 1) * C::kTileBytes);
          for (int t = 0; t < (it.has_tile1 ? 2 : 1); ++t)
            for (int c = 0; c < D / 64; ++c)
              ptx::tma_load_4d(sbase + C::kSmemQ + t * C::kTileBytes + c * C::kSubBytes, mq, B.q_full,
                               c * 64, it.head, qrow + t * kTileM, it.bcoord);
          int gg = g;
          for (int j = 0; j < it.nmax; ++j, ++gg) {
            const int s = gg % C::kStages;
            const uint32_t ph = (gg / C::kStages) & 1;
            if (gg >= C::kStages) ptx::mbar_wait(B.k_empty + 8 * s, ph ^ 1);
            ptx::mbar_arrive_expect_tx(B.k_full + 8 * s, C::kTileBytes);
            for (int c = 0; c < D / 64; ++c)
              ptx::tma_load_4d(sbase + C::kSmemK + s * C::kTileBytes + c * C::kSubBytes, mk,
                               B.k_full + 8 * s, c * 64, it.hk, krow + j * kTileN, it.bcoord);
            if (gg >= C::kStages) ptx::mbar_wait(B.v_empty + 8 * s, ph ^ 1);
            ptx::mbar_arrive_expect_tx(B.v_full + 8 * s, C::kTileBytes);
            for (int c = 0; c < D / 64; ++c)
              ptx::tma_load_4d(sbase + C::kSmemV + s * C::kTileBytes + c * C::kSubBytes, mv,
                               B.v_full + 8 * s, c * 64, it.hk, krow + j * kTileN, it.bcoord);
          }
        }
        ++nq;
        g += it.nmax;
      }
    } else if (warp == kMmaWarp && ptx::elect_one()) {
      // ---- MMA issuer ------------------------------------------------------------
      constexpr uint32_t fmt = kBF16 ? ptx::kFmtBF16 : ptx::kFmtF16;
      constexpr uint32_t idesc_qk = ptx::make_idesc(fmt, ptx::kMajorK, ptx::kMajorK, kTileM, kTileN);
      constexpr uint32_t idesc_pv = ptx::make_idesc(fmt, ptx::kMajorK, ptx::kMajorMN, kTileM, D);
      constexpr uint64_t hi_k = ptx::make_smem_desc_hi(16, 1024);             // K-major operands
      constexpr uint64_t hi_mn = ptx::make_smem_desc_hi(C::kSubBytes, 1024);  // V: MN-major
      // descriptors are rebuilt each iteration from a laundered base (ptx::smem_desc_at): a hoisted
      // table of 64-bit descriptors does not fit this warp's 40 registers
      uint32_t b16 = sbase >> 4;
      auto qk = [&](int t, int stage) {
        const uint32_t ka16 = b16 + stage * (C::kTileBytes >> 4);
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          const uint32_t off = (ks / 4) * C::kSubBytes + (ks % 4) * 32;
          ptx::mma_ss(tmem + t * 128, ptx::smem_desc_at(hi_k, b16 + t * (C::kTileBytes >> 4), C::kSmemQ + off),
                      ptx::smem_desc_at(hi_k, ka16, C::kSmemK + off), idesc_qk | a.qk_negate, ks > 0);
        }
      };
      // O_t (+)= P_t V: the two 64-key halves of P arrive separately (p_lo, p_hi)
      auto pv = [&](int t, int stage, bool acc, uint32_t parity) {
        const uint32_t va16 = b16 + stage * (C::kTileBytes >> 4);
#pragma unroll
        for (int ks = 0; ks < kTileN / 16; ++ks) {
          if (ks == 0 || ks == kTileN / 32) {
            ptx::mbar_wait((ks ? B.p_hi : B.p_lo) + 8 * t, parity);
            ptx::tc_fence_after();
          }
          ptx::mma_ts(tmem + 256 + t * 128, tmem + t * 128 + kPOff + ks * 8,
                      ptx::smem_desc_at(hi_mn, va16, C::kSmemV + ks * 2048), idesc_pv, acc || ks > 0);
        }
      };
      FA_TRACE_DECL(tr, 0, tr_cta);
      int nq = 0, g = 0;             // as in the TMA lane
      int cnt[2] = {0, 0};           // S tiles issued so far per query tile (s_full / p / pv_done phases)
      int used[2] = {0, 0};          // items that used O_t so far (o_free phases)
      for (int n = 0;; ++n) {
        const int L = next_item(n);
        if (L < 0) break;
        const Item it = decode(L);
        if (!it.valid || it.nmax == 0) continue;
        const int nt0 = it.ntile[0], nt1 = it.ntile[1], nmax = it.nmax;
        FA_TRACE_EV(tr, 1);
        ptx::mbar_wait(B.q_full, nq & 1);
        ++nq;
        ptx::tc_fence_after();
        ptx::mbar_wait(B.k_full + 8 * (g % C::kStages), (g / C::kStages) & 1);
        ptx::tc_fence_after();
        FA_TRACE_EV(tr, 20);
        b16 = ptx::opaque(sbase >> 4);
        if (nt0 > 0) { qk(0, g % C::kStages); ptx::mma_commit(B.s_full); }
        if (nt1 > 0) { qk(1, g % C::kStages); ptx::mma_commit(B.s_full + 8); }
        ptx::mma_commit(B.k_empty + 8 * (g % C::kStages));
        if (nmax == 1) ptx::mma_commit(B.q_empty);           // no further S = Q K^T in this item
        FA_TRACE_EV(tr, 21);
        for (int j = 0; j < nmax; ++j, ++g) {
          const int s = g % C::kStages, sn = (g + 1) % C::kStages;
          const uint32_t ph = (g / C::kStages) & 1, phn = ((g + 1) / C::kStages) & 1;
          b16 = ptx::opaque(sbase >> 4);
          FA_TRACE_EV(tr, 2);
          ptx::mbar_wait(B.v_full + 8 * s, ph);
          FA_TRACE_EV(tr, 3);
          if (j < nt0) {
            FA_TRACE_EV(tr, 4);
            // the first P V of an item overwrites O_0: the previous item's epilogue must have read it
            if (j == 0 && used[0] >= 1) { ptx::mbar_wait(B.o_free, (used[0] - 1) & 1); ptx::tc_fence_after(); }
            pv(0, s, j > 0, (cnt[0] + j) & 1);
            ptx::mma_commit(B.pv_done);
          }
          if (j + 1 < nmax) {
            ptx::mbar_wait(B.k_full + 8 * sn, phn);
            ptx::tc_fence_after();
          }
          FA_TRACE_EV(tr, 5);
          if (j + 1 < nt0) { qk(0, sn); ptx::mma_commit(B.s_full); }
          FA_TRACE_EV(tr, 6);
          if (j < nt1) {
            FA_TRACE_EV(tr, 7);
            if (j == 0 && used[1] >= 1) { ptx::mbar_wait(B.o_free + 8, (used[1] - 1) & 1); ptx::tc_fence_after(); }
            pv(1, s, j > 0, (cnt[1] + j) & 1);
            ptx::mma_commit(B.pv_done + 8);
          }
          ptx::mma_commit(B.v_empty + 8 * s);
          if (j + 1 < nt1) { qk(1, sn); ptx::mma_commit(B.s_full + 8); }
          if (j + 1 < nmax) ptx::mma_commit(B.k_empty + 8 * sn);
          if (j + 2 == nmax) ptx::mma_commit(B.q_empty);     // that was the item's last S = Q K^T
        }
        cnt[0] += nt0; cnt[1] += nt1;
        used[0] += nt0 > 0; used[1] += nt1 > 0;
      }
    }
  } else if ((warp >> 2) >= C::kQTiles) {
    ptx::setmaxnreg_dec<40>();                        // D = 256: the second softmax warpgroup has no tile
  } else {
    // ======================= softmax warpgroups ===================================
    ptx::setmaxnreg_inc<232>();
    const int t = warp >> 2;                          // query tile of this warpgroup
    const int r_in_tile = (warp & 3) * 32 + lane;     // TMEM lane == row inside the tile
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t tS = tmem + t * 128 + lane_sel;
    const uint32_t tP = tS + kPOff;
    const uint32_t tO = tmem + 256 + t * 128 + lane_sel;
    // with an additive bias the scores are moved to the scaled log2 domain when they are read
    // (s <- s * scale_log2 + bias * log2 e), and everything after works with a unit scale
    const float c2 = kMask ? 1.f : a.scale_log2;
    DropConst dc;
    if constexpr (kDropout) {
      dc.add7 = (0x80u - (a.drop_threshold & 0x7fu)) * 0x01010101u;
      dc.t_msb = a.drop_threshold >= 128u;
    }
    int cnt = 0;                                      // S tiles of my query tile consumed so far
    FA_TRACE_DECL(trs, 1 + t, tr_cta && (warp & 3) == 0 && lane == 0);
    for (int n = 0;; ++n) {
    const int L = next_item(n);
    if (L < 0) break;
    const Item it = decode(L);
    if (!it.valid) continue;
    const int bidx = it.bidx, head = it.head, sq = it.sq, sk = it.sk, shift = it.shift, m0 = it.m0, q0 = it.q0;
    const int row = m0 + t * kTileM + r_in_tile;      // row inside the sequence
    const int n_t = it.ntile[t];
    const uint16_t *mrow = nullptr;
    if (kMask && row < sq)
      mrow = static_cast<const uint16_t *>(a.mask) +
             (((size_t)(a.mask_b == 1 ? 0 : bidx) * a.mask_h + (a.mask_h == 1 ? 0 : head)) * a.mask_sq + row) * (size_t)a.mask_sk;
    const size_t msk = kMask ? (size_t)a.mask_sk : 0;         // row stride of the bias (== sk for dense inputs)
    const bool mvec = (msk % 8 == 0) && ((reinterpret_cast<uintptr_t>(a.mask) & 15) == 0);
    const uint16_t *mbase = kMask ? static_cast<const uint16_t *>(a.mask) +
        ((size_t)(a.mask_b == 1 ? 0 : bidx) * a.mask_h + (a.mask_h == 1 ? 0 : head)) * (size_t)a.mask_sq * msk : nullptr;
    const int row_limit = row + shift;                // last visible key when causal
    const int tile_row0 = m0 + t * kTileM;
    uint32_t bh = 0;
    if constexpr (kDropout) bh = (uint32_t)((a.batch_base + bidx) * a.h + head);

    float m_run = -INFINITY, l_run = 0.f;
    // Dropout (csrc/philox_spec.h: one Philox call per 16 columns, one byte per column).  The verdicts do
    // not depend on S, so they are drawn AHEAD, in the shadow of the MUFU stream: while one 64-column half is
    // exponentiated, the loop steps the four Philox calls of the half that is handed over next -- (j, 1)
    // during half 0, (j + 1, 0) during half 1 -- five rounds per 8 exponentials, round-robin over the
    // calls (four independent chains).  keepw[16 h + w]: bit 7 of each byte = keep, half h of the tile.
    uint32_t keepw[kDropout ? 32 : 1];
    uint32_t pc[4][4];
    auto ctr_init = [&](int q, int j, int hf) {
      pc[q][0] = (uint32_t)row; pc[q][1] = (uint32_t)(j * (kTileN / 16) + hf * 4 + q); pc[q][2] = bh; pc[q][3] = a.ctr3;
    };
    if constexpr (kDropout) {
      FA_TRACE_EV(trs, 17);
      if (n_t > 0) {                      // the item's first half: whole calls, before S arrives
#pragma unroll
        for (int q = 0; q < 4; ++q) ctr_init(q, 0, 0);
#pragma unroll
        for (int r = 0; r < FA_PHILOX_ROUNDS; ++r)
#pragma unroll
          for (int q = 0; q < 4; ++q) fa_philox_round(pc[q], a.key0, a.key1, r);
#pragma unroll
        for (int w = 0; w < 16; ++w) keepw[w] = ptx::opaque(keep_word(pc[w / 4][w % 4], dc));
      }
    }
    for (int j = 0; j < n_t; ++j) {
      // attn_mask, whole tiles of 16-byte aligned rows (non-dropout builds, which own a 4 KB staging block per
      // warp): a lane per ROW would fetch 16 bytes from each of 32 cache lines per instruction.  Instead the warp
      // reads its 32 rows x 128 columns as whole 128-byte lines (four rows per instruction), before the wait
      // for S -- the bias does not depend on it -- and turns them into rows through the staging block below.
      // The first 64 columns are parked in the block right away, the other 64 wait in registers.
      uint4 mpre[kMaskStage ? 8 : 1];
      bool mfast = false;
      if constexpr (kMaskStage) {
        mfast = mvec && (j * kTileN + kTileN <= sk);
        if (mfast) {
          const uint16_t *mtile = mbase + (size_t)(tile_row0 + (warp & 3) * 32) * msk + j * kTileN;
          const int rows_in = sq - (tile_row0 + (warp & 3) * 32);
          const uint32_t mstage = sbase + C::kSmemO + (uint32_t)(t * 4 + (warp & 3)) * 4096u;
          uint4 m0v[8];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int r = (i % 8) * 4 + (lane >> 3);
            const uint4 v = r < rows_in ? __ldg(reinterpret_cast<const uint4 *>(mtile + (size_t)r * msk + (i / 8) * 64 + (lane & 7) * 8))
                                        : make_uint4(0, 0, 0, 0);
            if (i < 8) m0v[i] = v; else mpre[i - 8] = v;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = i * 4 + (lane >> 3), ch = lane & 7;
            st_shared_v4(mstage + r * 128 + ((ch ^ (r & 7)) << 4), m0v[i].x, m0v[i].y, m0v[i].z, m0v[i].w);
          }
        }
      }
      FA_TRACE_EV(trs, 10);
      ptx::mbar_wait(B.s_full + 8 * t, (cnt + j) & 1);
      ptx::tc_fence_after();
      FA_TRACE_EV(trs, 11);
      uint32_t s[128];
      {
        uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
        uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
        uint32_t(&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
        uint32_t(&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
        ptx::tmem_ld_x32(tS, s0);
        ptx::tmem_ld_x32(tS + 32, s1);
        ptx::tmem_ld_x32(tS + 64, s2);
        ptx::tmem_ld_x32(tS + 96, s3);
        ptx::tmem_wait_ld();
      }
      FA_TRACE_EV(trs, 12);
      const int col0 = j * kTileN;
      if (kMaskStage && mfast) {                               // uniform: attn_mask through the staging block
        const float sc = a.scale_log2;
        const uint32_t mstage = sbase + C::kSmemO + (uint32_t)(t * 4 + (warp & 3)) * 4096u;
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          if (hf == 1) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int r = i * 4 + (lane >> 3), ch = lane & 7;
              st_shared_v4(mstage + r * 128 + ((ch ^ (r & 7)) << 4), mpre[i].x, mpre[i].y, mpre[i].z, mpre[i].w);
            }
          }
          __syncwarp();
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const uint4 v = ld_shared_v4(mstage + lane * 128 + ((k ^ (lane & 7)) << 4));
            const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const uint16_t mb = (uint16_t)(w4[e / 2] >> (16 * (e & 1)));
              const float bias = kBF16 ? __uint_as_float((uint32_t)mb << 16)
                                       : __half2float(*reinterpret_cast<const __half *>(&mb));
              const int c = hf * 64 + k * 8 + e;
              s[c] = __float_as_uint(fmaf(__uint_as_float(s[c]), sc, bias * kLog2e));
            }
          }
          __syncwarp();
        }
      } else if constexpr (kMask) {                            // uniform: attn_mask (reference flash_attn.h:35-36)
        const float sc = a.scale_log2;
#pragma unroll
        for (int c8 = 0; c8 < kTileN; c8 += 8) {
          uint16_t mb[8];
          const int col = col0 + c8;
          if (mrow && mvec && col + 8 <= sk) {
            *reinterpret_cast<uint4 *>(mb) = *reinterpret_cast<const uint4 *>(mrow + col);
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) mb[e] = (mrow && col + e < sk) ? mrow[col + e] : (uint16_t)0;
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const float bias = kBF16 ? __uint_as_float((uint32_t)mb[e] << 16)
                                     : __half2float(*reinterpret_cast<const __half *>(&mb[e]));
            s[c8 + e] = __float_as_uint(fmaf(__uint_as_float(s[c8 + e]), sc, bias * kLog2e));
          }
        }
      }
      // masking is needed only on the diagonal and on the ragged last tile (warp-uniform test)
      const bool need_mask = (col0 + kTileN > sk) || (a.causal && col0 + kTileN - 1 > tile_row0 + shift);
      if (need_mask) {
        const int lim = a.causal ? min(sk - 1, row_limit) : sk - 1;
#pragma unroll
        for (int c = 0; c < 128; ++c)
          if (col0 + c > lim) s[c] = 0xff800000u;  // -inf
      }
      float mx0 = m_run, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int c = 0; c < 128; c += 8) {
        mx0 = ptx::max3(mx0, __uint_as_float(s[c]), __uint_as_float(s[c + 1]));
        mx1 = ptx::max3(mx1, __uint_as_float(s[c + 2]), __uint_as_float(s[c + 3]));
        mx2 = ptx::max3(mx2, __uint_as_float(s[c + 4]), __uint_as_float(s[c + 5]));
        mx3 = ptx::max3(mx3, __uint_as_float(s[c + 6]), __uint_as_float(s[c + 7]));
      }
      const float m_new = fmaxf(ptx::max3(mx0, mx1, mx2), mx3);
      if (j == 0) {
        m_run = m_new;
      } else {
        // lazy rescale: only when some row of this warp grew by more than 2^kRescaleThreshold
        const bool grow = (m_new - m_run) * c2 > kRescaleThreshold || (m_run == -INFINITY && m_new > m_run);
        if (__any_sync(0xffffffffu, grow)) {
          const float alpha = m_new == -INFINITY ? 1.f : ptx::ex2((m_run - m_new) * c2);
          l_run *= alpha;
          m_run = m_new;
          ptx::mbar_wait(B.pv_done + 8 * t, (cnt + j - 1) & 1);   // O_t holds tiles 0..j-1
          ptx::tc_fence_after();
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            ptx::tmem_ld_x32(tO + c * 32, o);
            ptx::tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            ptx::tmem_st_x32(tO + c * 32, o);
          }
          ptx::tmem_wait_st();
        }
      }
      FA_TRACE_EV(trs, 13);
      const float m_use = m_run == -INFINITY ? 0.f : m_run;
      const float neg_mc = -m_use * c2;
      const float2 c2v = make_float2(c2, c2), nmv = make_float2(neg_mc, neg_mc);
      float2 l01 = make_float2(0.f, 0.f), l23 = make_float2(0.f, 0.f);
      // P = exp2(S c2 - m c2) for 64 columns -> 32 packed words, in place (pairs land below the
      // read cursor).  kEmu of every 8 exponentials run on the FMA pipe (ptx::ex2_poly2), the rest
      // on MUFU; tiles with masked (-inf) entries use MUFU only so that masked P is exactly 0.
      auto exp_half = [&](auto emu_tag, int half) {
        constexpr int kEmu = decltype(emu_tag)::value;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
#pragma unroll
          for (int pr = 0; pr < 4; ++pr) {
            const int c = half * 64 + g * 8 + 2 * pr;
            float2 x = __ffma2_rn(ptx::u2f2(s[c], s[c + 1]), c2v, nmv);
            if (2 * pr < kEmu) {
              x = ptx::ex2_poly2(x);
            } else {
              x.x = ptx::ex2(x.x);
              x.y = ptx::ex2(x.y);
            }
            if (pr & 1) l23 = __fadd2_rn(l23, x);
            else l01 = __fadd2_rn(l01, x);
            s[c / 2] = pack_pair(x.x, x.y, kBF16);
          }
          if constexpr (kDropout) {
            if (kApplyInGroup) {            // this group's 8 entries: keep words 2g, 2g + 1 of the current half
#pragma unroll
              for (int u = 0; u < 2; ++u) {
                uint32_t m01, m23;
                expand_keep(keepw[half * 16 + 2 * g + u], m01, m23);
                s[half * 32 + 4 * g + 2 * u] &= m01;
                s[half * 32 + 4 * g + 2 * u + 1] &= m23;
              }
            }
#pragma unroll
            for (int k = kPhiloxSteps * g / 8; k < kPhiloxSteps * (g + 1) / 8; ++k)
              fa_philox_round(pc[k % 4], a.key0, a.key1, k / 4);
            // nvvm would sink the rounds to their use; ptxas would sink them below the tile's last
            // exponential.  Laundering pins them to this group, __syncwarp() ends ptxas's scheduling block.
#pragma unroll
            for (int q = 0; q < 4; ++q)
#pragma unroll
              for (int i = 0; i < 4; ++i) pc[q][i] = ptx::opaque(pc[q][i]);
            if (kSyncGroups && g % kSyncGroups == kSyncGroups - 1) __syncwarp();
          }
        }
      };
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        if constexpr (kDropout) {
#pragma unroll
          for (int q = 0; q < 4; ++q) ctr_init(q, half ? j + 1 : j, 1 - half);
        }
        // D = 64 without dropout is the most exponential-bound build (half the MMA work per exp2): there 4 of
        // every 8 exponentials on the FMA pipe pay (635 -> 690 TFLOP/s at S4096 causal, 679 -> 780 without the mask,
        // varlen mix 480 -> 520-580; host switch a.use_emu: off below 1536 keys); at D = 128 they gain
        // 4 % at S4096 but lose 2 % at S2048, and with dropout they wreck the draw-ahead schedule (881 -> 629)
        constexpr int kEmu = kMask ? 0 : (kEmuPer8 == 0 && D == 64 && !kDropout) ? 4 : kEmuPer8;   // a bias may hold -inf: ex2_poly2 would give 2^-125, not 0
        if (kEmu == 0 || need_mask || !a.use_emu) exp_half(std::integral_constant<int, 0>{}, half);
        else exp_half(std::integral_constant<int, kEmu>{}, half);
        if constexpr (kDropout) {
          if (!kApplyInGroup) {
#pragma unroll
            for (int w = half * 16; w < half * 16 + 16; ++w) {
              uint32_t m01, m23;
              expand_keep(keepw[w], m01, m23);
              s[w * 2] &= m01;
              s[w * 2 + 1] &= m23;
            }
          }
#pragma unroll
          for (int w = 0; w < 16; ++w) keepw[(1 - half) * 16 + w] = keep_word(pc[w / 4][w % 4], dc);
        }
        // hand P over in halves: O += P[:, :64] V[:64] starts while the second half is exponentiated
        uint32_t(&ph)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[half * 32]);
        ptx::tmem_st_x32(tP + half * 32, ph);
        ptx::tmem_wait_st();
        ptx::tc_fence_before();
        ptx::mbar_arrive((half ? B.p_hi : B.p_lo) + 8 * t);
        if (half == 0) FA_TRACE_EV(trs, 14);
      }
      l_run += (l01.x + l01.y) + (l23.x + l23.y);
      FA_TRACE_EV(trs, 15);
    }

    // ---- epilogue: O_t / l -> global, LSE ------------------------------------------
    const bool row_ok = row < sq;
    if constexpr (kStageO) {
      // A lane owns one ROW of O; writing it out directly costs one 16-byte store per lane and row (32 cache
      // lines per instruction: ~5000 cycles per tile, traced).  Each warp instead stages 64 columns of its 32
      // rows in its own 4 KB of shared memory (16-byte chunks XOR-swizzled by row) and writes them back
      // as whole 128-byte lines, four rows per instruction.
      const uint32_t stage = sbase + C::kSmemO + (uint32_t)(t * 4 + (warp & 3)) * 4096u;
      const int wrow0 = tile_row0 + (warp & 3) * 32;                                           // the warp's first row
      uint16_t *wbase = static_cast<uint16_t *>(a.out) + (a.cu_q ? (size_t)q0 * a.o_row : (size_t)bidx * a.o_batch) +
                        (size_t)wrow0 * a.o_row + (size_t)head * a.o_head;
      const int rows_left = sq - (tile_row0 + (warp & 3) * 32);                                // rows of this warp inside the sequence
      const size_t row_pitch = (size_t)a.o_row;
      if (n_t > 0) {
        ptx::mbar_wait(B.pv_done + 8 * t, (cnt + n_t - 1) & 1);
        ptx::tc_fence_after();
      }
      const float inv = (n_t > 0 && l_run > 0.f) ? a.rp_dropout / l_run : 0.f;
  #pragma unroll
      for (int hf = 0; hf < D / 64; ++hf) {
        if (hf * 64 >= a.d) break;
  #pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t o[32];
          if (n_t > 0) {
            ptx::tmem_ld_x32(tO + hf * 64 + c * 32, o);
            ptx::tmem_wait_ld();
          } else {
  #pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = 0;
          }
  #pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int ch = c * 4 + i;
            st_shared_v4(stage + lane * 128 + ((ch ^ (lane & 7)) << 4),
                         pack_pair(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv, kBF16),
                         pack_pair(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv, kBF16),
                         pack_pair(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv, kBF16),
                         pack_pair(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv, kBF16));
          }
        }
        if (hf == D / 64 - 1 || (hf + 1) * 64 >= a.d) {
          if (n_t > 0) {
            ptx::tc_fence_before();
            ptx::mbar_arrive(B.o_free + 8 * t);            // O_t may be overwritten by the next item
          }
        }
        __syncwarp();
  #pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = i * 4 + (lane >> 3), ch = lane & 7;
          const uint4 v = ld_shared_v4(stage + r * 128 + ((ch ^ (r & 7)) << 4));
          if (r < rows_left && hf * 64 + ch * 8 < a.d)
            *reinterpret_cast<uint4 *>(wbase + (size_t)r * row_pitch + hf * 64 + ch * 8) = v;
        }
        __syncwarp();
      }
    } else {
      // direct: one 16-byte store per lane and row (the dropout builds, see kStageO)
      const size_t tok = (size_t)(a.cu_q ? q0 : bidx * a.sq) + row;
      uint4 *orow = reinterpret_cast<uint4 *>(static_cast<uint16_t *>(a.out) + (tok * a.h + head) * a.d);
      if (n_t > 0) {
        ptx::mbar_wait(B.pv_done + 8 * t, (cnt + n_t - 1) & 1);
        ptx::tc_fence_after();
        const float inv = l_run > 0.f ? a.rp_dropout / l_run : 0.f;
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
          uint32_t o[32];
          ptx::tmem_ld_x32(tO + c * 32, o);
          ptx::tmem_wait_ld();
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 v;
              v.x = pack_pair(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv, kBF16);
              v.y = pack_pair(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv, kBF16);
              v.z = pack_pair(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv, kBF16);
              v.w = pack_pair(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv, kBF16);
              if ((c * 4 + i) * 8 < a.d) orow[c * 4 + i] = v;
            }
          }
        }
        ptx::tc_fence_before();
        ptx::mbar_arrive(B.o_free + 8 * t);              // O_t may be overwritten by the next item
      } else if (row_ok) {
#pragma unroll
        for (int i = 0; i < D / 8; ++i)
          if (i * 8 < a.d) orow[i] = make_uint4(0, 0, 0, 0);
      }
    }
    if (row_ok) {
      const float m_use = m_run == -INFINITY ? 0.f : m_run;
      a.lse[((size_t)bidx * a.h + head) * a.sq + row] =
          l_run > 0.f ? (log2f(l_run) + m_use * c2) * kLn2 : INFINITY;
    }
    FA_TRACE_EV(trs, 16);
    cnt += n_t;
    }   // items
  }

  // ---- teardown ----------------------------------------------------------------------
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 512);
  }
}

}  // namespace

// ---- host side ---------------------------------------------------------------------

int trace_read_fwd(unsigned long long *dst, int role, int max_events) {
#ifdef FA_TRACE
  return trace::read_events(dst, role, max_events);
#else
  (void)dst; (void)role; (void)max_events;
  return 0;
#endif
}

bool tc_fwd_supported(const Params &p) {
  if (p.d <= 0 || p.d > 256 || p.d % 8) return false;   // d is padded to 64, 128 or 256 by TMA zero fill
  if (!isfinite(p.scale)) return false;          // (negative: the MMA's negate-A bit; zero: see launch_tc_fwd_t)
  if (p.sq <= 0 || p.sk <= 0) return false;
  return true;
}

namespace {

std::atomic<SchedSlot *> g_fwd_sched_addr[kMaxCachedDevices];   // device address of g_fwd_sched, per device
std::atomic<unsigned> g_fwd_ticket{0};

template <int D, bool kBF16, bool kDropout, bool kMask>
const char *launch_tc_fwd_t(const Params &p, cudaStream_t stream) {
  using C = Cfg<D>;
  CUtensorMap tq, tk, tv;
  const bool packed = p.cu_q != nullptr;
  // dense: (d, heads, seqlen, batch); packed: (d, heads, total tokens, 1)
  // packed: the row count is a placeholder; each CTA bounds its copy of the map (ptx::patch_rows)
  const uint64_t rows_q = (uint64_t)p.sq, rows_k = (uint64_t)p.sk;
  const uint64_t nb = packed ? 1 : (uint64_t)p.b;
  const char *e;
  // Decode-like calls (one query row, GQA, no dropout / bias): the query heads of a K/V head's group become
  // the ROWS of one tile -- Q [b, 1, h, d] read as [b, group, hk, d] through a strided tensor map -- so a
  // group streams its K/V head once instead of once per query head.  LSE [b, h, 1] is the same memory
  // under either view; O goes out through the staged epilogue's strides.  With sq == 1 the (bottom-right
  // aligned) causal mask hides nothing, so the packed problem is never causal.
  const int group = p.hk > 0 ? p.h / p.hk : 1;
  const bool gqa_pack = !packed && p.sq == 1 && group > 1 && p.h % p.hk == 0 && !kDropout && !kMask;
  const int sq_eff = gqa_pack ? group : p.sq, h_eff = gqa_pack ? p.hk : p.h;
  if (gqa_pack) {
    if ((e = make_tmap_strided(&tq, p.q, kBF16, p.d, (uint64_t)p.hk, (uint64_t)group, (uint64_t)p.b,
                               (uint64_t)group * p.d, (uint64_t)p.d, (uint64_t)p.h * p.d, kTileM))) return e;
  } else if ((e = make_tmap_rows(&tq, p.q, kBF16, p.d, p.h, rows_q, nb, kTileM))) return e;
  if ((e = make_tmap_rows(&tk, p.k, kBF16, p.d, p.hk, rows_k, nb, kTileN))) return e;
  if ((e = make_tmap_rows(&tv, p.v, kBF16, p.d, p.hk, rows_k, nb, kTileN))) return e;

  TcFwdArgs a;
  a.out = p.out; a.lse = p.lse; a.cu_q = p.cu_q; a.cu_k = p.cu_k;
  a.b = p.b; a.sq = sq_eff; a.sk = p.sk; a.h = h_eff; a.hk = p.hk; a.batch_base = p.batch_base;
  a.o_batch = gqa_pack ? (long long)p.h * p.d : (long long)p.sq * p.h * p.d;
  a.o_row = gqa_pack ? (long long)p.d : (long long)p.h * p.d;
  a.o_head = gqa_pack ? (long long)group * p.d : (long long)p.d;
  a.d = p.d;
  a.mask = p.mask; a.mask_b = (int)p.mask_b; a.mask_h = (int)p.mask_h; a.mask_sq = (int)p.mask_sq; a.mask_sk = (int)p.mask_sk;
  // The running max is tracked on RAW scores, so the scale the kernel multiplies by must be positive: a negative
  // softmax_scale negates the scores instead (negate-A bit of the Q K^T MMA); softmax_scale == 0 (every visible key
  // weighs the same) runs with 1e-30: |s| * 1e-30 is below half an ulp of any fp32 exponent argument, so every
  // exponential is exactly 1 and LSE = log(visible keys), as with 0 -- while a hidden key's -inf stays -inf (0 * -inf
  // would be NaN).
  a.scale_log2 = (p.scale == 0.f ? 1e-30f : fabsf(p.scale)) * kLog2e;
  a.qk_negate = p.scale < 0.f ? (1u << 13) : 0u;
  a.rp_dropout = p.rp_dropout;
  a.drop_threshold = p.drop_threshold;
  a.key0 = (uint32_t)p.seed ^ (uint32_t)(p.offset >> 32);
  a.key1 = (uint32_t)(p.seed >> 32);
  a.ctr3 = (uint32_t)p.offset;
  a.causal = (p.causal && !gqa_pack) ? 1 : 0;
  a.use_emu = p.sk >= 1536 ? 1 : 0;

  auto kern = tc_fwd_kernel<D, kBF16, kDropout, kMask>;
  int dev = 0, sm_count = 0;
  if ((e = current_device(&dev, &sm_count))) return e;
  static std::atomic<uint64_t> smem_opt_in{0};   // per instantiation, one bit per device
  if ((e = ensure_dynamic_smem(kern, (int)C::smem_bytes(stage_o(kDropout)), dev, smem_opt_in))) return e;
  a.nqb = (sq_eff + C::kQTiles * kTileM - 1) / (C::kQTiles * kTileM);
  a.n_items = a.nqb * h_eff * p.b;
  // no more items than SMs: one item per CTA, no counter
  a.persistent = a.n_items > sm_count ? 1 : 0;
  a.sched = nullptr;
  int grid = a.n_items;
  if (a.persistent) {
    grid = sm_count;
    if ((e = take_sched_slot(g_fwd_sched, dev, g_fwd_sched_addr, g_fwd_ticket, &a.sched))) return e;
  }
  kern<<<grid, kThreads, C::smem_bytes(stage_o(kDropout)), stream>>>(tq, tk, tv, a);
  const cudaError_t err = cudaGetLastError();
  return err == cudaSuccess ? nullptr : cudaGetErrorString(err);
}

}  // namespace

const char *launch_tc_fwd(const Params &p, cudaStream_t stream) {
  const bool drop = p.p_dropout > 0.f;
#define FA_MASKED(DD, BF, DR) (p.mask ? launch_tc_fwd_t<DD, BF, DR, true>(p, stream) : launch_tc_fwd_t<DD, BF, DR, false>(p, stream))
#define FA_DISPATCH(DD)                                                                        \
  if (p.is_bf16) return drop ? FA_MASKED(DD, true, true)                        \
                             : FA_MASKED(DD, true, false);                      \
  return drop ? FA_MASKED(DD, false, true) : FA_MASKED(DD, false, false);
  if (p.d > 128) { FA_DISPATCH(256) }
  if (p.d > 64) { FA_DISPATCH(128) }
  FA_DISPATCH(64)
#undef FA_DISPATCH
#undef FA_MASKED
}

}  // namespace fa
