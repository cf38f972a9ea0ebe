// umma_probe.cu -- single-CTA bring-up probe for the tcgen05 building blocks the
// attention kernels rely on.  It runs, with descriptor fields taken from the
// command line so one GPU session can try several encodings:
//   (1) S = Q K^T : A, B from shared memory (TMA, 128B swizzle, K-major)
//   (2) O = P V   : A from tensor memory (bf16 written by tcgen05.st),
//                   B from shared memory (TMA, 128B swizzle, MN-major)
// and checks both against a CPU matmul.  Exit code 0 = both match.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 \
//        -I flash_attention_hip_b200/csrc tests/cuda/umma_probe.cu -o tests/cuda/umma_probe
#include <cuda_bf16.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ptx_sm100.cuh"
#include "tma_host.cuh"

using namespace fa;

struct ProbeArgs {
  int d;                 // head dim: 64 or 128
  uint32_t k_lbo, k_sbo; // K-major descriptor fields (bytes)
  uint32_t k_step;       // bytes added to the start address per 16-element K step
  uint32_t sub_stride;   // bytes between the two 64-column sub-tiles of a 128-wide operand
  uint32_t mn_lbo, mn_sbo, mn_step;
  uint32_t idesc_qk, idesc_pv;
  uint32_t p_col;        // TMEM column where P is staged
};

constexpr int kRows = 128;

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tk,
             const __grid_constant__ CUtensorMap tv, const float *__restrict__ p_in,
             float *__restrict__ s_out, float *__restrict__ o_out, ProbeArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  const uint32_t sbase = (ptx::smem_u32(smem) + 1023u) & ~1023u;
  const uint32_t tile_bytes = kRows * a.d * 2;
  const uint32_t sQ = sbase, sK = sbase + tile_bytes, sV = sbase + 2 * tile_bytes;
  const uint32_t bar_tma = ptx::smem_u32(&bars[0]), bar_mma = ptx::smem_u32(&bars[1]);

  if (tid == 0) {
    ptx::mbar_init(bar_tma, 1);
    ptx::mbar_init(bar_mma, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(ptx::smem_u32(&tmem_base_s), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const uint32_t tS = tmem, tO = tmem + 256, tP = tmem + a.p_col;

  if (tid == 0) {
    ptx::mbar_arrive_expect_tx(bar_tma, 3 * tile_bytes);
    for (int c = 0; c < a.d / 64; ++c) {
      ptx::tma_load_4d(sQ + c * a.sub_stride, &tq, bar_tma, c * 64, 0, 0, 0);
      ptx::tma_load_4d(sK + c * a.sub_stride, &tk, bar_tma, c * 64, 0, 0, 0);
      ptx::tma_load_4d(sV + c * a.sub_stride, &tv, bar_tma, c * 64, 0, 0, 0);
    }
    ptx::mbar_wait(bar_tma, 0);
    ptx::tc_fence_after();
    const uint64_t hi = ptx::make_smem_desc_hi(a.k_lbo, a.k_sbo);
    for (int ks = 0; ks < a.d / 16; ++ks) {
      const uint32_t off = (ks / 4) * a.sub_stride + (ks % 4) * a.k_step;
      ptx::mma_ss(tS, ptx::smem_desc(hi, sQ + off), ptx::smem_desc(hi, sK + off), a.idesc_qk, ks > 0);
    }
    ptx::mma_commit(bar_mma);
  }
  __syncwarp();
  ptx::mbar_wait(bar_mma, 0);
  ptx::tc_fence_after();

  const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
  {
    uint32_t r[32];
    for (int c = 0; c < 4; ++c) {
      ptx::tmem_ld_x32(tS + lane_base + c * 32, r);
      ptx::tmem_wait_ld();
      for (int i = 0; i < 32; ++i) s_out[tid * 128 + c * 32 + i] = __uint_as_float(r[i]);
    }
  }
  // stage P (bf16 pairs, element 2c in the low half of column c)
  {
    uint32_t r[32];
    for (int c = 0; c < 2; ++c) {
      for (int i = 0; i < 32; ++i) {
        const float lo = p_in[tid * 128 + c * 64 + 2 * i], hi = p_in[tid * 128 + c * 64 + 2 * i + 1];
        const __nv_bfloat162 pk = __floats2bfloat162_rn(lo, hi);
        r[i] = *reinterpret_cast<const uint32_t *>(&pk);
      }
      ptx::tmem_st_x32(tP + lane_base + c * 32, r);
    }
    ptx::tmem_wait_st();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (tid == 0) {
    ptx::tc_fence_after();
    const uint64_t hi = ptx::make_smem_desc_hi(a.mn_lbo, a.mn_sbo);
    for (int ks = 0; ks < kRows / 16; ++ks)
      ptx::mma_ts(tO, tP + ks * 8, ptx::smem_desc(hi, sV + ks * a.mn_step), a.idesc_pv, ks > 0);
    ptx::mma_commit(bar_mma);
  }
  __syncwarp();
  ptx::mbar_wait(bar_mma, 1);
  ptx::tc_fence_after();
  {
    uint32_t r[32];
    for (int c = 0; c < a.d / 32; ++c) {
      ptx::tmem_ld_x32(tO + lane_base + c * 32, r);
      ptx::tmem_wait_ld();
      for (int i = 0; i < 32; ++i) o_out[tid * a.d + c * 32 + i] = __uint_as_float(r[i]);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(tmem, 512);
}

static float bf16r(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 3; } } while (0)

int main(int argc, char **argv) {
  ProbeArgs a;
  a.d = 128;
  a.k_lbo = 16; a.k_sbo = 1024; a.k_step = 32; a.sub_stride = 0;
  a.mn_lbo = 0; a.mn_sbo = 1024; a.mn_step = 2048;
  a.idesc_qk = 0; a.idesc_pv = 0; a.p_col = 32;
  uint32_t b_major_pv = ptx::kMajorMN;
  for (int i = 1; i < argc; ++i) {
    char key[64]; unsigned long val;
    if (sscanf(argv[i], "%63[^=]=%lu", key, &val) != 2) { printf("bad arg %s\n", argv[i]); return 2; }
    if (!strcmp(key, "d")) a.d = (int)val;
    else if (!strcmp(key, "k_lbo")) a.k_lbo = val; else if (!strcmp(key, "k_sbo")) a.k_sbo = val;
    else if (!strcmp(key, "k_step")) a.k_step = val; else if (!strcmp(key, "sub_stride")) a.sub_stride = val;
    else if (!strcmp(key, "mn_lbo")) a.mn_lbo = val; else if (!strcmp(key, "mn_sbo")) a.mn_sbo = val;
    else if (!strcmp(key, "mn_step")) a.mn_step = val; else if (!strcmp(key, "p_col")) a.p_col = val;
    else if (!strcmp(key, "idesc_qk")) a.idesc_qk = val; else if (!strcmp(key, "idesc_pv")) a.idesc_pv = val;
    else { printf("unknown key %s\n", key); return 2; }
  }
  const int d = a.d;
  if (!a.sub_stride) a.sub_stride = kRows * 128;           // one [128 x 64] sub-tile
  if (!a.mn_lbo) a.mn_lbo = kRows * 128;                    // next 64-wide chunk of d
  if (!a.idesc_qk) a.idesc_qk = ptx::make_idesc(ptx::kFmtBF16, ptx::kMajorK, ptx::kMajorK, 128, 128);
  if (!a.idesc_pv) a.idesc_pv = ptx::make_idesc(ptx::kFmtBF16, ptx::kMajorK, b_major_pv, 128, d);

  std::vector<float> q(128 * d), k(128 * d), v(128 * d), p(128 * 128);
  srand(1);
  auto rnd = [] { return (rand() / (float)RAND_MAX) * 2.f - 1.f; };
  for (auto &x : q) x = bf16r(rnd());
  for (auto &x : k) x = bf16r(rnd());
  for (auto &x : v) x = bf16r(rnd());
  for (auto &x : p) x = bf16r(rand() / (float)RAND_MAX);
  std::vector<__nv_bfloat16> qb(128 * d), kb(128 * d), vb(128 * d);
  for (int i = 0; i < 128 * d; ++i) { qb[i] = __float2bfloat16_rn(q[i]); kb[i] = __float2bfloat16_rn(k[i]); vb[i] = __float2bfloat16_rn(v[i]); }

  __nv_bfloat16 *dq, *dk, *dv; float *dp, *ds, *dout;
  CK(cudaMalloc(&dq, 128 * d * 2)); CK(cudaMalloc(&dk, 128 * d * 2)); CK(cudaMalloc(&dv, 128 * d * 2));
  CK(cudaMalloc(&dp, 128 * 128 * 4)); CK(cudaMalloc(&ds, 128 * 128 * 4)); CK(cudaMalloc(&dout, 128 * d * 4));
  CK(cudaMemcpy(dq, qb.data(), 128 * d * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dk, kb.data(), 128 * d * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dv, vb.data(), 128 * d * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dp, p.data(), 128 * 128 * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(ds, 0, 128 * 128 * 4)); CK(cudaMemset(dout, 0, 128 * d * 4));

  CUtensorMap tq, tk, tv;
  const char *e;
  if ((e = make_tmap_rows(&tq, dq, true, d, 1, 128, 1, 128)) || (e = make_tmap_rows(&tk, dk, true, d, 1, 128, 1, 128)) ||
      (e = make_tmap_rows(&tv, dv, true, d, 1, 128, 1, 128))) { printf("tensor map: %s\n", e); return 3; }

  const int smem = 3 * 128 * d * 2 + 1024;
  CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  probe_kernel<<<1, 128, smem>>>(tq, tk, tv, dp, ds, dout, a);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());

  std::vector<float> s(128 * 128), o(128 * d);
  CK(cudaMemcpy(s.data(), ds, s.size() * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o.data(), dout, o.size() * 4, cudaMemcpyDeviceToHost));
  double es = 0, eo = 0;
  for (int i = 0; i < 128; ++i)
    for (int j = 0; j < 128; ++j) {
      double acc = 0;
      for (int c = 0; c < d; ++c) acc += (double)q[i * d + c] * k[j * d + c];
      es = fmax(es, fabs(acc - s[i * 128 + j]));
    }
  for (int i = 0; i < 128; ++i)
    for (int c = 0; c < d; ++c) {
      double acc = 0;
      for (int j = 0; j < 128; ++j) acc += (double)p[i * 128 + j] * v[j * d + c];
      eo = fmax(eo, fabs(acc - o[i * d + c]));
    }
  const bool oks = es < 1e-2, oko = eo < 1e-2;
  printf("PROBE d=%d k_lbo=%u k_sbo=%u k_step=%u mn_lbo=%u mn_sbo=%u mn_step=%u p_col=%u : qk %s (max err %.4g)  pv %s (max err %.4g)\n",
         d, a.k_lbo, a.k_sbo, a.k_step, a.mn_lbo, a.mn_sbo, a.mn_step, a.p_col, oks ? "PASS" : "FAIL", es,
         oko ? "PASS" : "FAIL", eo);
  return (oks && oko) ? 0 : 1;
}
