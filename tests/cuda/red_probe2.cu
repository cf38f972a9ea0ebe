// red_probe2.cu -- bulk reduce-add rate of ONE SM against block size and blocks in flight (developer probe).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int kDepth>
__global__ void __launch_bounds__(128, 1) probe(float *dst, int iters, int blk_bytes, size_t span_bytes) {
  extern __shared__ __align__(1024) uint8_t smem[];
  float *s = reinterpret_cast<float *>(smem);
  for (int i = threadIdx.x; i < kDepth * blk_bytes / 4; i += blockDim.x) s[i] = 1.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    size_t off = (size_t)blockIdx.x * (span_bytes / gridDim.x);
    const size_t lim = off + span_bytes / gridDim.x;
    const size_t base = off;
    for (int cc = 0; cc < iters; ++cc) {
      if (cc >= kDepth) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kDepth - 1) : "memory");
      const uint32_t buf = smem_u32(smem) + (cc % kDepth) * blk_bytes;
      asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                   ::"l"(reinterpret_cast<uint8_t *>(dst) + off), "r"(buf), "r"(blk_bytes) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      off += blk_bytes;
      if (off >= lim) off = base;
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}
template <int kDepth>
void run(float *dst, int ctas, int blk, size_t span) {
  const int iters = (64 << 20) / blk;   // 64 MB per CTA
  cudaFuncSetAttribute(probe<kDepth>, cudaFuncAttributeMaxDynamicSharedMemorySize, kDepth * blk + 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<kDepth><<<ctas, 128, kDepth * blk + 1024>>>(dst, 64, blk, span);
  cudaEventRecord(e0);
  probe<kDepth><<<ctas, 128, kDepth * blk + 1024>>>(dst, iters, blk, span);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("ctas %3d  block %5d B  in flight %d: %7.1f GB/s per CTA  %.2f TB/s total (%s)\n", ctas, blk, kDepth,
         (double)iters * blk / ms / 1e6, (double)ctas * iters * blk / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  const size_t span = 64ull << 20;
  float *dst; cudaMalloc(&dst, span); cudaMemset(dst, 0, span);
  for (int ctas : {1, 148})
    for (int blk : {4096, 8192, 16384}) {
      run<2>(dst, ctas, blk, span); run<4>(dst, ctas, blk, span); run<8>(dst, ctas, blk, span);
    }
  return 0;
}
