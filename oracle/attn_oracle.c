/*
 * attn_oracle.c -- CPU oracle for the flash_attn.h hot path.  TEST
 * INFRASTRUCTURE: see the header comment of attn_oracle.h (who may load it,
 * and why parity is "unpinned by the reference").
 *
 * Build: make -C oracle   ->  oracle/_build/libattn_oracle.so
 */
#include "attn_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../flash_attention_hip_b200/csrc/philox_spec.h"

/* ---- storage conversions (IEEE round-to-nearest-even) ------------------- */

static inline uint32_t f32_bits(float x) { uint32_t u; memcpy(&u, &x, 4); return u; }
static inline float bits_f32(uint32_t u) { float x; memcpy(&x, &u, 4); return x; }

static uint16_t f32_to_bf16(float x) {
  uint32_t u = f32_bits(x);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u); /* quiet NaN */
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

static float bf16_to_f32(uint16_t h) { return bits_f32((uint32_t)h << 16); }

static uint16_t f32_to_f16(float x) {
  const uint32_t u = f32_bits(x);
  const uint32_t sign = (u >> 16) & 0x8000u;
  const uint32_t a = u & 0x7fffffffu;
  if (a > 0x7f800000u) return (uint16_t)(sign | 0x7e00u);            /* NaN */
  if (a >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);           /* >= 65520 -> inf */
  if (a < 0x33000001u) return (uint16_t)sign;                        /* < 2^-25 -> 0 */
  int32_t e = (int32_t)(a >> 23) - 127;
  uint32_t m = (a & 0x7fffffu) | 0x800000u;
  uint32_t shift, half;
  if (e < -14) { shift = (uint32_t)(13 + (-14 - e)); }               /* subnormal half */
  else { shift = 13; }
  half = m >> shift;
  const uint32_t rem = m & ((1u << shift) - 1u);
  const uint32_t halfway = 1u << (shift - 1);
  if (rem > halfway || (rem == halfway && (half & 1u))) half++;
  if (e < -14) return (uint16_t)(sign | half);                        /* may carry into normal */
  half += (uint32_t)(e + 15 - 1) << 10;                               /* implicit bit adds 1<<10 */
  return (uint16_t)(sign | half);
}

static float f16_to_f32(uint16_t h) {
  const uint32_t sign = ((uint32_t)h & 0x8000u) << 16;
  const uint32_t e = (h >> 10) & 0x1fu;
  const uint32_t m = h & 0x3ffu;
  if (e == 0) {
    if (m == 0) return bits_f32(sign);
    const float f = (float)m * (1.0f / 16777216.0f); /* m * 2^-24 */
    return sign ? -f : f;
  }
  if (e == 31) return bits_f32(sign | 0x7f800000u | (m << 13));
  return bits_f32(sign | ((e + 112u) << 23) | (m << 13));
}

uint16_t fa_oracle_f32_to_storage(float x, int is_bf16) {
  return is_bf16 ? f32_to_bf16(x) : f32_to_f16(x);
}

float fa_oracle_storage_to_f32(uint16_t x, int is_bf16) {
  return is_bf16 ? bf16_to_f32(x) : f16_to_f32(x);
}

/* ---- the two precision builds ------------------------------------------ */

#define REAL float
#define SUFFIX(n) n##_f32
#define EXP(x) expf(x)
#define LOG(x) logf(x)
#include "attn_oracle_impl.inc"
#undef REAL
#undef SUFFIX
#undef EXP
#undef LOG

#define REAL double
#define SUFFIX(n) n##_f64
#define EXP(x) exp(x)
#define LOG(x) log(x)
#include "attn_oracle_impl.inc"
#undef REAL
#undef SUFFIX
#undef EXP
#undef LOG

/* ---- public entry points ------------------------------------------------ */

static int check(const fa_oracle_params *p) {
  if (!p || p->batch < 0 || p->num_heads <= 0 || p->num_heads_k <= 0 || p->head_size <= 0) return 1;
  if (p->num_heads % p->num_heads_k) return 1;
  if ((p->cu_seqlens_q == NULL) != (p->cu_seqlens_k == NULL)) return 1;
  return 0;
}

int fa_oracle_fwd(const fa_oracle_params *p, const uint16_t *q, const uint16_t *k,
                  const uint16_t *v, uint16_t *out, float *lse, float *out_f32,
                  uint16_t *softmax_ret) {
  if (check(p)) return 2;
  return p->use_fp64 ? fwd_f64(p, q, k, v, out, lse, out_f32, softmax_ret)
                     : fwd_f32(p, q, k, v, out, lse, out_f32, softmax_ret);
}

int fa_oracle_bwd(const fa_oracle_params *p, const uint16_t *dout, const uint16_t *q,
                  const uint16_t *k, const uint16_t *v, const uint16_t *out,
                  const float *lse, uint16_t *dq, uint16_t *dk, uint16_t *dv,
                  float *dq_f32, float *dk_f32, float *dv_f32) {
  if (check(p)) return 2;
  return p->use_fp64
             ? bwd_f64(p, dout, q, k, v, out, lse, dq, dk, dv, dq_f32, dk_f32, dv_f32)
             : bwd_f32(p, dout, q, k, v, out, lse, dq, dk, dv, dq_f32, dk_f32, dv_f32);
}

void fa_oracle_dropout_mask(uint64_t seed, uint64_t offset, float p_dropout, uint32_t bh,
                            int32_t sq, int32_t sk, uint8_t *mask) {
  const uint32_t thr = fa_dropout_threshold(p_dropout);
  for (int32_t i = 0; i < sq; ++i)
    for (int32_t j = 0; j < sk; ++j)
      mask[(size_t)i * sk + j] =
          (uint8_t)fa_dropout_keep(seed, offset, bh, (uint32_t)i, (uint32_t)j, thr);
}

void fa_oracle_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  const fa_philox4_t r = fa_philox4x32(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
  for (int i = 0; i < 4; ++i) out[i] = r.v[i];
}

void fa_oracle_philox_r(const uint32_t ctr[4], const uint32_t key[2], int rounds, uint32_t out[4]) {
  const fa_philox4_t r = fa_philox4x32_r(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1], rounds);
  for (int i = 0; i < 4; ++i) out[i] = r.v[i];
}

int fa_oracle_philox_rounds(void) { return FA_PHILOX_ROUNDS; }

int fa_oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
