// red_probe.cu -- how fast can B200 take fp32 bulk reduce-adds (cp.reduce.async.bulk ... add.f32) from all SMs?
// The tcgen05 backward sends every dQ tile (128 x 128 fp32 = 64 KB per (query tile, key tile) pair) to the
// caller's dq_accum this way; this probe issues the same 16 KB blocks with the same double-buffered
// staging and nothing else, so its rate is the ceiling that traffic puts on the backward:
//   flops per pair = 5 GEMMs x 2 x 128^3 = 20.97 MFLOP per 65536 B  ->  320 flop per reduced byte.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo red_probe.cu -o red_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <bool kReduce>
__global__ void __launch_bounds__(128, 1) probe(float *dst, int iters, int tiles_per_cta, size_t span_tiles) {
  extern __shared__ __align__(1024) uint8_t smem[];
  float *s = reinterpret_cast<float *>(smem);
  for (int i = threadIdx.x; i < 2 * 4096; i += blockDim.x) s[i] = 1.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    int cc = 0;
    for (int it = 0; it < iters; ++it) {
      // one "dQ tile" = 4 blocks of 16 KB; CTAs walk different tiles, wrapping inside span_tiles
      const size_t tile = ((size_t)blockIdx.x * tiles_per_cta + it) % span_tiles;
      float *g = dst + tile * 16384;
      for (int c = 0; c < 4; ++c, ++cc) {
        if (cc >= 2) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
        const uint32_t buf = smem_u32(smem) + (cc & 1) * 16384;
        if (kReduce)
          asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                       ::"l"(g + c * 4096), "r"(buf), "r"(16384) : "memory");
        else
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                       ::"l"(g + c * 4096), "r"(buf), "r"(16384) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int iters = 2048;
  // footprints: 64 MB (fits the 126 MB L2) and 1 GB (the C3 backward's dq_accum: spills to HBM)
  for (size_t span_mb : {64, 1024}) {
    const size_t span_tiles = span_mb * 1024 * 1024 / 65536;
    float *dst;
    cudaMalloc(&dst, span_tiles * 65536);
    cudaMemset(dst, 0, span_tiles * 65536);
    for (int red = 1; red >= 0; --red) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      auto k = red ? probe<true> : probe<false>;
      cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 + 1024);
      k<<<sms, 128, 32768 + 1024>>>(dst, 64, 1024, span_tiles);
      cudaEventRecord(e0);
      k<<<sms, 128, 32768 + 1024>>>(dst, iters, (int)(span_tiles / sms ? span_tiles / sms : 1), span_tiles);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double bytes = (double)sms * iters * 65536.0;
      const cudaError_t err = cudaGetLastError();
      printf("%s  footprint %4zu MB  %d CTAs x %d tiles of 64 KB: %.3f ms  %.2f TB/s  -> backward ceiling %.0f TFLOP/s (%s)\n",
             red ? "bulk reduce-add f32" : "bulk store         ", span_mb, sms, iters, ms, bytes / ms / 1e9,
             bytes / ms / 1e9 * 320.0, cudaGetErrorString(err));
    }
    cudaFree(dst);
  }
  // per-SM or chip-wide limit?  Same blocks from fewer CTAs (64 MB footprint).
  {
    const size_t span_tiles = 64ull * 1024 * 1024 / 65536;
    float *dst;
    cudaMalloc(&dst, span_tiles * 65536);
    cudaMemset(dst, 0, span_tiles * 65536);
    for (int ctas : {1, 8, 37, 74, 148}) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaFuncSetAttribute(probe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 + 1024);
      cudaEventRecord(e0);
      probe<true><<<ctas, 128, 32768 + 1024>>>(dst, iters, (int)(span_tiles / ctas), span_tiles);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      const double bytes = (double)ctas * iters * 65536.0;
      printf("bulk reduce-add f32  %3d CTAs: %.3f ms  %.2f TB/s total  %.1f GB/s per CTA\n", ctas, ms, bytes / ms / 1e9,
             bytes / ms / 1e6 / ctas);
    }
    cudaFree(dst);
  }
  return 0;
}
