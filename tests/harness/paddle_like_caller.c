/*
 * paddle_like_caller.c -- a framework-shaped caller of the flash_attn.h ABI.
 *
 * Plain C.  It does what Paddle's dynamic-loader table does with the reference library
 * (/root/reference/flash_attn.h:10-138): dlopen() the .so, bind the six symbols BY NAME, allocate
 * the scratch buffers with the sizes the header implies (softmax_d fp32 [b, h, sq_rounded],
 * dq_accum fp32 [b, sq_rounded, h, d_rounded]), run flash_attn_fwd -> flash_attn_bwd reusing the
 * forward's seed/offset, and check the bool / flash_attn_error() convention.  Inputs come from a
 * small LCG that tests/test_gpu_harness.py restates in numpy; outputs go to a binary file the test
 * compares with the oracle.
 *
 *   paddle_like_caller <libflash_attn.so> <out.bin> [b sq sk h hk d causal p_dropout]
 */
#include <cuda_runtime_api.h>
#include <dlfcn.h>
#include <math.h>
#include <stdbool.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* The reference prototypes, restated for C (include/flash_attn.h is the maintained copy). */
typedef struct CUstream_st *stream_t;
typedef bool (*fwd_fn)(const void *, const void *, const void *, void *, void *, void *, void *, int, int, int,
                       int, int, int, int, int, int, float, float, float, bool, bool, bool, stream_t, uint64_t,
                       uint64_t, const void *, const int64_t *);
typedef bool (*bwd_fn)(const void *, const void *, const void *, const void *, const void *, const void *,
                       const void *, void *, void *, void *, void *, void *, int, int, int, int, int, int, int,
                       int, int, float, float, float, bool, bool, int, stream_t, uint64_t, uint64_t,
                       const void *, const int64_t *);
typedef void (*set_error_fn)(const char *);
typedef const char *(*error_fn)(void);

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_));                     \
      return 2;                                                                    \
    }                                                                              \
  } while (0)

static uint32_t lcg_state;
/* bf16 bits of a value in (-1, 1) with 7 mantissa bits: exactly representable, so the numpy
 * restatement produces the same bits without any rounding question. */
static uint16_t next_bf16(void) {
  lcg_state = lcg_state * 1664525u + 1013904223u;
  const int v = (int)((lcg_state >> 16) & 0xff) - 128;  /* -128..127 */
  const float f = (float)v / 128.0f;
  uint32_t u;
  memcpy(&u, &f, 4);
  return (uint16_t)(u >> 16);
}

static int round_up(int x, int m) { return (x + m - 1) / m * m; }

int main(int argc, char **argv) {
  if (argc < 3) {
    fprintf(stderr, "usage: %s <libflash_attn.so> <out.bin> [b sq sk h hk d causal p_dropout]\n", argv[0]);
    return 2;
  }
  int b = 2, sq = 256, sk = 256, h = 4, hk = 2, d = 128, causal = 1;
  float p_dropout = 0.1f;
  if (argc >= 11) {
    b = atoi(argv[3]); sq = atoi(argv[4]); sk = atoi(argv[5]); h = atoi(argv[6]); hk = atoi(argv[7]);
    d = atoi(argv[8]); causal = atoi(argv[9]); p_dropout = (float)atof(argv[10]);
  }
  void *lib = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
  if (!lib) {
    fprintf(stderr, "dlopen: %s\n", dlerror());
    return 2;
  }
  fwd_fn fwd = (fwd_fn)dlsym(lib, "flash_attn_fwd");
  bwd_fn bwd = (bwd_fn)dlsym(lib, "flash_attn_bwd");
  void *vfwd = dlsym(lib, "flash_attn_varlen_fwd"), *vbwd = dlsym(lib, "flash_attn_varlen_bwd");
  set_error_fn set_error = (set_error_fn)dlsym(lib, "flash_attn_set_error");
  error_fn error = (error_fn)dlsym(lib, "flash_attn_error");
  if (!fwd || !bwd || !vfwd || !vbwd || !set_error || !error) {
    fprintf(stderr, "a reference symbol is missing from %s\n", argv[1]);
    return 2;
  }

  const size_t nq = (size_t)b * sq * h * d, nk = (size_t)b * sk * hk * d;
  const int sq_r = round_up(sq, 128), sk_r = round_up(sk, 128), d_r = round_up(d, 32);
  uint16_t *hq = malloc(nq * 2), *hk_ = malloc(nk * 2), *hv = malloc(nk * 2), *hdo = malloc(nq * 2);
  lcg_state = 12345u;
  for (size_t i = 0; i < nq; ++i) hq[i] = next_bf16();
  for (size_t i = 0; i < nk; ++i) hk_[i] = next_bf16();
  for (size_t i = 0; i < nk; ++i) hv[i] = next_bf16();
  for (size_t i = 0; i < nq; ++i) hdo[i] = next_bf16();

  void *q, *k, *v, *dout, *out, *lse, *rng, *dq, *dk, *dv, *softmax_d, *dq_accum;
  cudaStream_t stream;
  CK(cudaStreamCreate(&stream));
  CK(cudaMalloc(&q, nq * 2)); CK(cudaMalloc(&dout, nq * 2)); CK(cudaMalloc(&out, nq * 2)); CK(cudaMalloc(&dq, nq * 2));
  CK(cudaMalloc(&k, nk * 2)); CK(cudaMalloc(&v, nk * 2)); CK(cudaMalloc(&dk, nk * 2)); CK(cudaMalloc(&dv, nk * 2));
  CK(cudaMalloc(&lse, (size_t)b * h * sq * 4)); CK(cudaMalloc(&rng, 16));
  CK(cudaMalloc(&softmax_d, (size_t)b * h * sq_r * 4));
  CK(cudaMalloc(&dq_accum, (size_t)b * sq_r * h * d_r * 4));
  CK(cudaMemcpyAsync(q, hq, nq * 2, cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(k, hk_, nk * 2, cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(v, hv, nk * 2, cudaMemcpyHostToDevice, stream));
  CK(cudaMemcpyAsync(dout, hdo, nq * 2, cudaMemcpyHostToDevice, stream));

  const float scale = 1.0f / sqrtf((float)d);
  const uint64_t seed = 1234, offset = 0;

  /* 1. the error convention: a bad call returns false and leaves a message */
  set_error("");
  if (fwd(q, k, v, rng, out, NULL, lse, b, sq, sk, sq_r, sk_r, h, /*num_heads_k=*/3, d, d_r, 0.f, scale, 1.f / scale,
          causal, false, true, stream, seed, offset, NULL, NULL)) {
    fprintf(stderr, "num_heads %% num_heads_k != 0 was accepted\n");
    return 1;
  }
  if (strlen(error()) == 0) {
    fprintf(stderr, "failure left no message\n");
    return 1;
  }
  printf("expected failure: %s\n", error());

  /* 2. forward then backward, as a training step would */
  if (!fwd(q, k, v, rng, out, NULL, lse, b, sq, sk, sq_r, sk_r, h, hk, d, d_r, p_dropout, scale, 1.f / scale, causal,
           false, true, stream, seed, offset, NULL, NULL)) {
    fprintf(stderr, "flash_attn_fwd: %s\n", error());
    return 1;
  }
  uint64_t hrng[2];
  CK(cudaMemcpyAsync(hrng, rng, 16, cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));
  if (!bwd(dout, q, k, v, out, softmax_d, lse, rng, dq, dk, dv, dq_accum, b, sq, sk, sq_r, sk_r, h, hk, d, d_r,
           p_dropout, scale, 1.f / scale, causal, true, /*num_splits=*/0, stream, hrng[0], hrng[1], NULL, NULL)) {
    fprintf(stderr, "flash_attn_bwd: %s\n", error());
    return 1;
  }

  uint16_t *ho = malloc(nq * 2), *hdq = malloc(nq * 2), *hdk = malloc(nk * 2), *hdv = malloc(nk * 2);
  float *hlse = malloc((size_t)b * h * sq * 4);
  CK(cudaMemcpyAsync(ho, out, nq * 2, cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(hdq, dq, nq * 2, cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(hdk, dk, nk * 2, cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(hdv, dv, nk * 2, cudaMemcpyDeviceToHost, stream));
  CK(cudaMemcpyAsync(hlse, lse, (size_t)b * h * sq * 4, cudaMemcpyDeviceToHost, stream));
  CK(cudaStreamSynchronize(stream));

  FILE *f = fopen(argv[2], "wb");
  if (!f) return 2;
  const int32_t hdr[8] = {b, sq, sk, h, hk, d, causal, (int32_t)(hrng[0])};
  fwrite(hdr, 4, 8, f);
  fwrite(ho, 2, nq, f); fwrite(hlse, 4, (size_t)b * h * sq, f);
  fwrite(hdq, 2, nq, f); fwrite(hdk, 2, nk, f); fwrite(hdv, 2, nk, f);
  fclose(f);
  printf("ok: fwd+bwd through dlsym-bound symbols, rng_state = {%llu, %llu}\n", (unsigned long long)hrng[0],
         (unsigned long long)hrng[1]);
  dlclose(lib);
  return 0;
}
