// pipe_probe.cu -- per-SMSP issue cost of the instructions the softmax loops are made of (B200).
// One CTA of W warps per SMSP x 4 SMSPs; each warp runs kIters x 8 independent instructions of one
// kind; reports cycles per warp-instruction per SMSP.   nvcc -arch=sm_100a -O3 -o pipe_probe pipe_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

constexpr int kIters = 512;

template <int OP>
__global__ void probe(float *out, long long *cycles, float seed) {
  float x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = seed + threadIdx.x * 1e-3f + i;
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i])); }
      if (OP == 1) { asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(x[i])); }
      if (OP == 2) {   // packed f32x2 FMA
        if (i < 4) {
          unsigned long long v;
          asm volatile("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(x[2 * i]), "f"(x[2 * i + 1]));
          asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(v));
          asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(x[2 * i]), "=f"(x[2 * i + 1]) : "l"(v));
        }
      }
      if (OP == 3) {   // F2FP pack
        uint32_t p;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p) : "f"(x[i]), "f"(x[(i + 1) & 7]));
        acc ^= p;
      }
      if (OP == 4) { asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(x[(i + 1) & 7]), "f"(x[(i + 2) & 7])); }
      if (OP == 5) {   // IMAD.WIDE
        unsigned long long v;
        asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(v) : "r"(__float_as_uint(x[i])), "r"(0xD2511F53u));
        x[i] = __uint_as_float((uint32_t)(v >> 32) ^ (uint32_t)v);
      }
      if (OP == 6) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(acc) : "r"(__float_as_uint(x[i])), "r"(acc + i)); }
      if (OP == 7) { asm volatile("add.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(x[(i + 1) & 7])); }
    }
  }
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char *name, int per_iter) {
  float *out; long long *cyc;
  cudaMalloc(&out, 1024 * 4); cudaMalloc(&cyc, 8);
  for (int warps_per_smsp = 1; warps_per_smsp <= 4; warps_per_smsp *= 2) {
    const int threads = warps_per_smsp * 4 * 32;
    probe<OP><<<1, threads>>>(out, cyc, 0.5f);
    probe<OP><<<1, threads>>>(out, cyc, 0.5f);
    long long c;
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    const double per = (double)c / ((double)kIters * per_iter * warps_per_smsp);
    printf("%-28s warps/SMSP=%d  %.2f cycles per warp-instruction per SMSP\n", name, warps_per_smsp, per);
  }
}

int main() {
  run<0>("MUFU.EX2", 8);
  run<1>("FFMA", 8);
  run<2>("FFMA2 (f32x2)", 4);
  run<3>("F2FP.BF16.PACK_AB", 8);
  run<4>("FMNMX3", 8);
  run<5>("IMAD.WIDE (+LOP)", 8);
  run<6>("LOP3", 8);
  run<7>("FADD", 8);
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
