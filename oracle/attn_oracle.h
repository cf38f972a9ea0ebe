/*
 * attn_oracle.h -- CPU oracle for the flash_attn.h hot path.  TEST INFRASTRUCTURE.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this.  The product library (libflash_attn.so) never
 * links, loads or calls it.
 *
 * PARITY UNPINNED BY THE REFERENCE: /root/reference holds only flash_attn.h
 * (no implementation, no tests, no golden vectors; SURVEY.md section 0 and
 * 8(c)), so there is nothing of the reference's to compile into oracle/_ref
 * and no reference fixture to check against.  This oracle is the "fp32 CPU
 * reference of the reference's semantics" that BASELINE.json's north_star
 * prescribes.  It is pinned three ways instead (tests/test_oracle.py):
 *   - the Philox core against the published Random123 known-answer vectors;
 *   - fp32 vs its own fp64 build vs an independent numpy float64 restatement;
 *   - backward against central finite differences of the fp64 forward.
 * Its outputs on fixed seeds are committed under tests/golden/.
 *
 * Layouts and argument meaning follow /root/reference/flash_attn.h:
 *   dense  q [b, sq, h, d], k/v [b, sk, h_k, d]            (:10-12)
 *   packed q [total_q, h, d], k/v [total_k, h_k, d]        (:38-40)
 *   cu_seqlens int32 [b+1]                                 (:41-42)
 *   GQA: query head h reads kv head h / (h / h_k)          (:22-23)
 *   lse fp32 [b, h, seqlen_q] (max_seqlen_q when packed)   (:74, :107)
 *   p_dropout = probability to drop                        (:89)
 * Everything the header leaves open follows DESIGN.md section 3 and
 * csrc/philox_spec.h.
 */
#ifndef FA_B200_ATTN_ORACLE_H_
#define FA_B200_ATTN_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
  int32_t batch;
  int32_t seqlen_q;      /* dense: the length; packed: max_seqlen_q (lse stride) */
  int32_t seqlen_k;
  int32_t num_heads;
  int32_t num_heads_k;
  int32_t head_size;
  int32_t is_causal;
  int32_t is_bf16;       /* 1 = bf16 storage, 0 = fp16 storage */
  int32_t use_fp64;      /* 1 = accumulate in double (self-validation build) */
  int32_t seqlen_q_rounded;   /* strides of softmax_ret only */
  int32_t seqlen_k_rounded;
  float p_dropout;
  float softmax_scale;
  uint64_t seed;
  uint64_t offset;
  const int32_t *cu_seqlens_q;   /* NULL = dense */
  const int32_t *cu_seqlens_k;
  const uint16_t *attn_mask;     /* additive bias, storage dtype, or NULL */
  int64_t mask_dims[4];          /* {b|1, h|1, sq, sk} */
  int32_t batch_base;            /* global index of batch 0 (dropout stream of a batch shard) */
  int32_t round_intermediates;   /* 1 = also round P (after dropout) and dS to the storage dtype before their
                                    GEMMs, as ANY 16-bit tensor-core pipeline must: the "bf16-rounded reference"
                                    whose error against the exact build calibrates the parity tests' second
                                    criterion (kernel error <= 2x this build's error).  0 = exact (the oracle). */
} fa_oracle_params;

/* Forward.  out: storage dtype, same layout as q.  lse: fp32 [b,h,seqlen_q].
 * out_f32 (optional): un-rounded output.  softmax_ret (optional): storage
 * dtype [b,h,sq_rounded,sk_rounded], probabilities, negated where dropped.
 * Returns 0 on success. */
int fa_oracle_fwd(const fa_oracle_params *p, const uint16_t *q, const uint16_t *k,
                  const uint16_t *v, uint16_t *out, float *lse, float *out_f32,
                  uint16_t *softmax_ret);

/* Backward from (dout, q, k, v, out, lse).  dq/dk/dv in storage dtype;
 * dq_f32/dk_f32/dv_f32 (optional) un-rounded. */
int fa_oracle_bwd(const fa_oracle_params *p, const uint16_t *dout, const uint16_t *q,
                  const uint16_t *k, const uint16_t *v, const uint16_t *out,
                  const float *lse, uint16_t *dq, uint16_t *dk, uint16_t *dv,
                  float *dq_f32, float *dk_f32, float *dv_f32);

/* Dropout keep mask for one (b, h): mask[sq*sk] bytes, 1 = keep. */
void fa_oracle_dropout_mask(uint64_t seed, uint64_t offset, float p_dropout, uint32_t bh,
                            int32_t sq, int32_t sk, uint8_t *mask);

/* Raw Philox-4x32 with the specification's round count (csrc/philox_spec.h: 7), and with an explicit
 * round count, for the known-answer tests (Random123 publishes vectors for 7 and 10 rounds). */
void fa_oracle_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void fa_oracle_philox_r(const uint32_t ctr[4], const uint32_t key[2], int rounds, uint32_t out[4]);
int fa_oracle_philox_rounds(void);

/* Storage conversions used by the oracle (round-to-nearest-even). */
uint16_t fa_oracle_f32_to_storage(float x, int is_bf16);
float fa_oracle_storage_to_f32(uint16_t x, int is_bf16);

int fa_oracle_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
