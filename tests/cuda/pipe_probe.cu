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

// Two instruction kinds on the SAME schedulers: warps 0-3 run A, warps 4-7 run B (one of each per
// SMSP).  If the pipes overlap, each side takes about as long as it does alone.
template <int OP>
__device__ __forceinline__ void pair_loop(float (&x)[8], uint32_t &acc) {
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i])); }
      if (OP == 5) {
        unsigned long long v;
        asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(v) : "r"(__float_as_uint(x[i])), "r"(0xD2511F53u));
        x[i] = __uint_as_float((uint32_t)(v >> 32) ^ (uint32_t)v);
      }
      if (OP == 6) { uint32_t u = __float_as_uint(x[i]); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u) : "r"(acc), "r"(i)); x[i] = __uint_as_float(u); }
      if (OP == 1) { asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(x[i])); }
    }
  }
}
template <int A, int B2>
__global__ void probe_pair(float *out, long long *cycles, float seed) {
  float x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = seed + threadIdx.x * 1e-3f + i;
  uint32_t acc = threadIdx.x;
  const bool first = threadIdx.x < 128;
  __syncthreads();
  const long long t0 = clock64();
  if (first) pair_loop<A>(x, acc);
  else pair_loop<B2>(x, acc);
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  out[threadIdx.x] = s + acc;
  if ((threadIdx.x & 127) == 0) cycles[threadIdx.x >> 7] = t1 - t0;
}

template <int A, int B2>
void run_pair(const char *name) {
  float *out; long long *cyc;
  cudaMalloc(&out, 1024 * 4); cudaMalloc(&cyc, 16);
  probe_pair<A, B2><<<1, 256>>>(out, cyc, 0.5f);
  probe_pair<A, B2><<<1, 256>>>(out, cyc, 0.5f);
  long long c[2];
  cudaMemcpy(c, cyc, 16, cudaMemcpyDeviceToHost);
  printf("%-28s together: A %.2f, B %.2f cycles per warp-instruction\n", name, (double)c[0] / (kIters * 8.0),
         (double)c[1] / (kIters * 8.0));
}

// Philox-4x32-10 throughput: kCalls independent calls per thread per iteration, W warps per SMSP.
__global__ void probe_philox(uint32_t *out, long long *cycles, uint32_t k0, uint32_t k1) {
  uint32_t acc = 0;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < 64; ++it) {
#pragma unroll
    for (int cb = 0; cb < 8; ++cb) {
      uint32_t c0 = threadIdx.x, c1 = it * 8 + cb, c2 = 77, c3 = 5, ka = k0, kb = k1;
#pragma unroll
      for (int r = 0; r < 10; ++r) {
        const unsigned long long p0 = (unsigned long long)0xD2511F53u * c0, p1 = (unsigned long long)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ ka, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ kb;
        c1 = (uint32_t)p1; c3 = (uint32_t)p0; c0 = n0; c2 = n2;
        ka += 0x9E3779B9u; kb += 0xBB67AE85u;
      }
      acc += (c0 ^ c1) + (c2 ^ c3);
    }
  }
  const long long t1 = clock64();
  out[threadIdx.x] = acc;
  if (threadIdx.x == 0) cycles[0] = t1 - t0;
}

void run_philox() {
  uint32_t *out; long long *cyc;
  cudaMalloc(&out, 1024 * 4); cudaMalloc(&cyc, 8);
  for (int w = 1; w <= 4; ++w) {
    probe_philox<<<1, w * 128>>>(out, cyc, 1, 2);
    probe_philox<<<1, w * 128>>>(out, cyc, 1, 2);
    long long c;
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("Philox-4x32-10 bare          warps/SMSP=%d  %.1f cycles per call (one warp) per SMSP\n", w, (double)c / (64.0 * 8 * w));
  }
}

int main() {
  run_pair<0, 5>("MUFU.EX2 | IMAD.WIDE(+LOP)");
  run_pair<0, 6>("MUFU.EX2 | LOP3");
  run_pair<5, 6>("IMAD.WIDE(+LOP) | LOP3");
  run_pair<1, 6>("FFMA | LOP3");
  run_pair<1, 5>("FFMA | IMAD.WIDE(+LOP)");
  run_pair<0, 0>("MUFU.EX2 | MUFU.EX2");
  run_philox();
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
