// ptx_sm100.cuh -- inline-PTX wrappers for the sm_100a features the attention
// kernels use: mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma /
// commit / ld / st / fences), UMMA shared-memory and instruction descriptors.
// Hand-written; field layouts follow the PTX ISA "tcgen05" chapter.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace fa {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------- mbarrier

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}

// Makes mbarrier.init visible to the async proxy (TMA, tcgen05.commit).
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
}

// Waits for the phase with the given parity to complete.  A wait that lasts
// longer than ~2^32 SM cycles is a protocol bug: trap instead of hanging the
// GPU (the launch then fails with an error the host reports).
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > (1ll << 32)) __trap();
  }
}

// ---------------------------------------------------------------------- TMA

__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap *m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}

// 4-D tiled load global -> shared, completion on an mbarrier (complete_tx).
__device__ __forceinline__ void tma_load_4d(uint32_t smem_dst, const CUtensorMap *m, uint32_t bar,
                                            int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// 4-D tiled store shared -> global (bulk group completion).
__device__ __forceinline__ void tma_store_4d(const CUtensorMap *m, uint32_t smem_src, int c0, int c1,
                                             int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
      ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ------------------------------------------------------------------ tcgen05

// One warp allocates `ncols` (power of two >= 32) TMEM columns; the base
// address is written to shared memory at `dst`.
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
               "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc]; one thread issues for the CTA.
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                       uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                       uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// mbarrier arrives (count 1) once every tcgen05.mma issued so far by this
// thread has completed.  Implies tcgen05.fence::before_thread_sync.
__device__ __forceinline__ void mma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}

// TMEM -> registers, shape 32x32b: lane i of the warp reads TMEM lane
// (lane_base + i), `N` consecutive 32-bit columns starting at taddr's column.
__device__ __forceinline__ void tmem_ld_x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}

__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// registers -> TMEM, shape 32x32b.x32
__device__ __forceinline__ void tmem_st_x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
        "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]),
        "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}

__device__ __forceinline__ void tmem_st_x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}

template <int N>
__device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

__device__ __forceinline__ void named_bar_arrive(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_sync(uint32_t id, uint32_t nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}



// ------------------------------------------------------------ fast math
// MUFU.EX2 without the denormal-range fix-up code exp2f() carries (3 extra
// instructions per call); inputs here are <= ~8, tiny results flush to zero.
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 3-input max (FMNMX3, sm_100+).
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float2 u2f2(uint32_t lo, uint32_t hi) {
  return make_float2(__uint_as_float(lo), __uint_as_float(hi));
}
// 2^x for a pair on the FMA pipe (no MUFU: B200's 16 ex2/clk/SM is what bounds a 128x128 softmax
// tile).  Cody-Waite: t = x + 1.5*2^23 leaves round(x) in t's low mantissa bits, xf = x - round(x)
// in [-0.5, 0.5], degree-3 minimax polynomial (max relative error 7.5e-5, far below bf16 rounding),
// exponent added with an integer shift-add.  x is clamped at -125, so -inf gives 2^-125, not 0:
// callers use ex2() on tiles that hold masked entries.
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
  const float kM = 12582912.f;
  x.x = fmaxf(x.x, -125.f);
  x.y = fmaxf(x.y, -125.f);
  const float2 t = __fadd2_rn(x, make_float2(kM, kM));
  const float2 xi = __fadd2_rn(t, make_float2(-kM, -kM));
  const float2 xf = __ffma2_rn(xi, make_float2(-1.f, -1.f), x);
  float2 p = __ffma2_rn(make_float2(0.055171459913253784f, 0.055171459913253784f), xf,
                        make_float2(0.2426108568906784f, 0.2426108568906784f));
  p = __ffma2_rn(p, xf, make_float2(0.6932609677314758f, 0.6932609677314758f));
  p = __ffma2_rn(p, xf, make_float2(0.9999281167984009f, 0.9999281167984009f));
  float2 r;
  r.x = __uint_as_float(__float_as_uint(p.x) + (__float_as_uint(t.x) << 23));
  r.y = __uint_as_float(__float_as_uint(p.y) + (__float_as_uint(t.y) << 23));
  return r;
}

// ---------------------------------------------- device-side tensor-map patching
// Packed (varlen) inputs: the host cannot know a sequence's last row (cu_seqlens
// lives on the device), so each CTA bounds its own copy of the tensor map:
// copy the 128-byte map to shared memory, replace one global dimension, publish
// it to a global slot with tensormap.cp_fenceproxy, acquire it for the TMA unit.

__device__ __forceinline__ void tensormap_replace_dim2_smem(uint32_t smem_desc, uint32_t new_dim) {
  asm volatile("tensormap.replace.tile.global_dim.shared::cta.b1024.b32 [%0], 2, %1;" ::"r"(smem_desc),
               "r"(new_dim)
               : "memory");
}

// Whole warp (.sync.aligned): shared-memory copy -> global slot, release at gpu scope.
__device__ __forceinline__ void tensormap_publish(const void *gmem_slot, uint32_t smem_desc) {
  asm volatile(
      "tensormap.cp_fenceproxy.global.shared::cta.tensormap::generic.release.gpu.sync.aligned [%0], [%1], 128;"
      ::"l"(reinterpret_cast<uint64_t>(gmem_slot)), "r"(smem_desc)
      : "memory");
}

__device__ __forceinline__ void tensormap_acquire(const void *gmem_slot) {
  asm volatile("fence.proxy.tensormap::generic.acquire.gpu [%0], 128;" ::"l"(reinterpret_cast<uint64_t>(gmem_slot))
               : "memory");
}

__device__ __forceinline__ uint32_t smid() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(r));
  return r;
}

// Per-SM slots for patched tensor maps.  Every kernel that uses them needs more
// than half of an SM's shared memory, so at most one such CTA is resident per
// SM and slot [smid] has a single owner at any time.
constexpr int kMaxSm = 1024;
constexpr int kMapsPerSm = 8;
struct alignas(128) TensorMapSlot { uint8_t bytes[128]; };

// Warp-collective: slot <- param map with its row dimension (dim 2) replaced.
__device__ __forceinline__ const CUtensorMap *patch_rows(const CUtensorMap &param_map, TensorMapSlot *slot,
                                                         uint32_t smem_desc, uint8_t *smem_desc_gen,
                                                         uint32_t rows, int lane) {
  reinterpret_cast<uint32_t *>(smem_desc_gen)[lane] = reinterpret_cast<const uint32_t *>(&param_map)[lane];
  __syncwarp();
  if (lane == 0) tensormap_replace_dim2_smem(smem_desc, rows);
  __syncwarp();
  tensormap_publish(slot, smem_desc);
  tensormap_acquire(slot);
  return reinterpret_cast<const CUtensorMap *>(slot);
}

// ------------------------------------------------------- UMMA descriptors

constexpr uint32_t kFmtF16 = 0, kFmtBF16 = 1;
constexpr uint32_t kMajorK = 0, kMajorMN = 1;

// Instruction descriptor, kind::f16, fp32 accumulate.
//  [4,6) D format (1 = f32)  [7,10) A format  [10,13) B format
//  [15] A major  [16] B major  [17,23) N >> 3  [24,29) M >> 4
__host__ __device__ constexpr uint32_t make_idesc(uint32_t fmt, uint32_t a_major, uint32_t b_major,
                                                  uint32_t M, uint32_t N) {
  return (1u << 4) | (fmt << 7) | (fmt << 10) | (a_major << 15) | (b_major << 16) | ((N >> 3) << 17) |
         ((M >> 4) << 24);
}

// Shared-memory matrix descriptor, 128-byte swizzle.
//  [0,14) address >> 4   [16,30) leading byte offset >> 4   [32,46) stride byte offset >> 4
//  [46,48) version = 1 (sm_100)   [61,64) layout: 2 = SWIZZLE_128B
// K-major operand (rows of 128 B, 8-row atoms): SBO = 1024 (next 8 rows), LBO unused (1).
// MN-major operand (k-rows of 128 B holding 64 MN elements): SBO = 1024 (next
// 8 k-rows), LBO = byte distance between consecutive 64-element MN chunks.
__host__ __device__ constexpr uint64_t make_smem_desc_hi(uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t((lbo_bytes >> 4) & 0x3fff) << 16) | (uint64_t((sbo_bytes >> 4) & 0x3fff) << 32) |
         (uint64_t(1) << 46) | (uint64_t(2) << 61);
}
__device__ __forceinline__ uint64_t smem_desc(uint64_t hi, uint32_t smem_addr) {
  return hi | uint64_t((smem_addr >> 4) & 0x3fff);
}

// Descriptor from a laundered base: `base16` = (CTA smem base >> 4) passed through opaque() once per
// loop iteration, `off` a compile-time byte offset.  One IADD per descriptor instead of a hoisted
// (and, in a 32-register MMA warp, spilled) table of 64-bit constants.
__device__ __forceinline__ uint64_t smem_desc_at(uint64_t hi, uint32_t base16, uint32_t off) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(base16 + (off >> 4) + (uint32_t)hi), "r"((uint32_t)(hi >> 32)));
  return d;
}
__device__ __forceinline__ uint32_t opaque(uint32_t x) {
  asm volatile("" : "+r"(x));
  return x;
}

}  // namespace ptx
}  // namespace fa
