// fa_bwd_dev.cuh -- device helpers shared by the tcgen05 backward kernels (fa_tc_bwd.cu: head sizes up to 128,
// fa_tc_bwd_wide.cu: 136..256): 16-bit packing, shared-memory vector access, bulk reduce-add, and the
// dropout keep-bit machinery of csrc/philox_spec.h on TRANSPOSED coordinates (TMEM lane = key row).
#pragma once

#include "fa_common.cuh"
#include "ptx_sm100.cuh"

namespace fa {
namespace bwd_dev {

__device__ __forceinline__ uint32_t pack2(float lo, float hi, bool bf16) {
  if (bf16) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
  }
  const __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<const uint32_t *>(&v);
}

__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

__device__ __forceinline__ float4 ld_shared_f32x4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// 16 KB shared -> global, element-wise fp32 add (async proxy, bulk group).
__device__ __forceinline__ void bulk_reduce_add_f32(float *gdst, uint32_t smem_src, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_src), "r"(bytes) : "memory");
}

// 16 keep bits (bit k: byte k of the 4 Philox words < T), csrc/philox_spec.h.
__device__ __forceinline__ uint32_t keep_bits16(const fa_philox4_t &r, uint32_t add7, bool t_msb) {
  uint32_t bits = 0;
#pragma unroll
  for (int w = 0; w < 4; ++w) {
    const uint32_t x = r.v[w];
    const uint32_t ge7 = (x & 0x7f7f7f7fu) + add7;
    const uint32_t keep = (t_msb ? ~(x & ge7) : ~(x | ge7)) & 0x80808080u;
    // gather bits 7,15,23,31 into a nibble
    const uint32_t nib = (keep * 0x00204081u) >> 28;
    bits |= nib << (4 * w);
  }
  return bits;
}

// AND-mask of packed pair c (entries 2c, 2c + 1 of 32) from a word of keep bits (bit e = keep entry e).
// wk[k] = bits << (7 - k) puts entry k + 8m at the top of byte m; PRMT's sign-replicate mode (selector bit
// 3) then expands two such bytes to 16 bits each: one PRMT + one AND per pair, 8 shifts per 32 entries.
__device__ __forceinline__ uint32_t pair_mask(const uint32_t (&wk)[8], int pair) {
  const int e = 2 * pair, k = e % 8, m = e / 8;
  const uint32_t sel = (8u + m) * 0x0011u + (12u + m) * 0x1100u;   // bytes {a.m, a.m, b.m, b.m}, sign-replicated
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(wk[k]), "r"(wk[k + 1]), "r"(sel));   // sel folds to an immediate
  return d;
}

// One butterfly stage of four 16x16 bit transposes (two per word) across the lanes of each
// half-warp: lanes whose bit kS differs exchange the off-diagonal kS-wide blocks.
template <int kS>
__device__ __forceinline__ void bit_transpose_stage(uint32_t &w01, uint32_t &w23, int lane) {
  constexpr uint32_t m = kS == 8 ? 0x00ff00ffu : kS == 4 ? 0x0f0f0f0fu : kS == 2 ? 0x33333333u : 0x55555555u;
  const bool lower = (lane & kS) == 0;
  const uint32_t km = lower ? m : ~m;
  const uint32_t y01 = __shfl_xor_sync(0xffffffffu, w01, kS);
  const uint32_t y23 = __shfl_xor_sync(0xffffffffu, w23, kS);
  const uint32_t z01 = lower ? (y01 << kS) : (y01 >> kS);
  const uint32_t z23 = lower ? (y23 << kS) : (y23 >> kS);
  w01 = (w01 & km) | (z01 & ~km);
  w23 = (w23 & km) | (z23 & ~km);
}


// One stage of two 16x16 bit transposes (the two half-words of ONE word), as above.
template <int kS>
__device__ __forceinline__ void bit_transpose_stage1(uint32_t &w01, int lane) {
  constexpr uint32_t m = kS == 8 ? 0x00ff00ffu : kS == 4 ? 0x0f0f0f0fu : kS == 2 ? 0x33333333u : 0x55555555u;
  const bool lower = (lane & kS) == 0;
  const uint32_t km = lower ? m : ~m;
  const uint32_t y01 = __shfl_xor_sync(0xffffffffu, w01, kS);
  const uint32_t z01 = lower ? (y01 << kS) : (y01 >> kS);
  w01 = (w01 & km) | (z01 & ~km);
}

}  // namespace bwd_dev
}  // namespace fa
