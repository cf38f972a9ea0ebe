/*
 * flash_attn.h -- C ABI of the B200-native FlashAttention-2 library.
 *
 * Drop-in boundary for yuguo-Jack/flash-attention-hip: the six symbols below
 * have the same names, argument order, argument types and return convention
 * as the reference header (/root/reference/flash_attn.h:10-138), so a caller
 * that binds them by name (Paddle's dynamic-loader table) links unchanged.
 *
 * Differences that do not touch the ABI:
 *   - valid C as well as C++ (<stdint.h>/<stdbool.h> instead of <cstdint>);
 *     `bool` is one byte in both languages on x86-64 SysV.
 *   - the stream argument is `fa_stream_t`, a pointer to an opaque struct.
 *     hipStream_t (reference :3,:32) and cudaStream_t are both such pointers;
 *     pass a cudaStream_t. NULL means the legacy default stream.
 *
 * Every pointer except `mask_dims` is caller-owned DEVICE memory. The library
 * allocates nothing that outlives a call, never synchronises the device and
 * enqueues all work on `stream`.
 *
 * Return: true on success. On failure false, and flash_attn_error() returns a
 * thread-local message. Success leaves the previous message untouched.
 *
 * Semantics the reference header leaves open are fixed in DESIGN.md section 3
 * ("decisions") and restated briefly beside each argument here.
 *
 * The six prototypes restate the interface of yuguo-Jack/flash-attention-hip's
 * flash_attn.h (BSD 3-Clause License, Copyright (c) 2024, yuguo): an ABI has to
 * match token for token.  Everything behind them is this repository's own.
 */
#ifndef FLASH_ATTN_B200_FLASH_ATTN_H_
#define FLASH_ATTN_B200_FLASH_ATTN_H_

#include <stdbool.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Opaque stream handle; layout-compatible with cudaStream_t / hipStream_t. */
struct CUstream_st;
typedef struct CUstream_st *fa_stream_t;

/*
 * Dense-batch forward. Replaces reference flash_attn.h:10-36 (27 arguments).
 *
 *   q    [batch, seqlen_q, num_heads,   head_size]  fp16/bf16
 *   k,v  [batch, seqlen_k, num_heads_k, head_size]
 *   out  [batch, seqlen_q, num_heads,   head_size]
 *   softmax_lse_ptr  fp32 [batch, num_heads, seqlen_q]: natural-log
 *        log-sum-exp of the scaled, masked, pre-dropout scores; +inf for a
 *        fully masked row (whose output row is 0).
 *   rng_state  2 x uint64 {seed, offset}, written when non-NULL.
 *   softmax_ptr  only when return_softmax: input dtype,
 *        [batch, num_heads, seqlen_q_rounded, seqlen_k_rounded], the
 *        normalised probabilities, 0 where the causal mask hides the entry,
 *        negated where dropout removed it; rows >= seqlen_q and columns >=
 *        seqlen_k of each plane are not written.
 *   *_rounded  used only as strides of softmax_ptr / scratch, never to index
 *        q, k, v, out.
 *   p_dropout  probability to DROP, in [0, 1].
 *   softmax_unscale  1/softmax_scale; used only with attn_mask.
 *   is_causal  bottom-right aligned: query i sees keys j <= i + seqlen_k - seqlen_q.
 *   attn_mask  optional additive bias in the input dtype with HOST dims
 *        mask_dims[4] = {b or 1, h or 1, seqlen_q, seqlen_k}; added to
 *        softmax_scale * q.k before the softmax. NULL = none.  On the packed
 *        entry points the bias is padded: {b or 1, h or 1, max_seqlen_q,
 *        max_seqlen_k}, entry (row, col) counted from each sequence's first token.
 */
bool flash_attn_fwd(const void * const q,
                    const void * const k,
                    const void * const v,
                    void * const rng_state,
                    void * const out,
                    void * const softmax_ptr,
                    void * const softmax_lse_ptr,
                    const int batch_size,
                    const int seqlen_q,
                    const int seqlen_k,
                    const int seqlen_q_rounded,
                    const int seqlen_k_rounded,
                    const int num_heads,
                    const int num_heads_k,
                    const int head_size,
                    const int head_size_rounded,
                    const float p_dropout,
                    const float softmax_scale,
                    const float softmax_unscale,
                    const bool is_causal,
                    const bool return_softmax,
                    const bool is_bf16,
                    fa_stream_t stream,
                    uint64_t seed,
                    uint64_t offset,
                    const void * const attn_mask,
                    const int64_t * const mask_dims);

/*
 * Packed variable-length forward. Replaces reference flash_attn.h:38-66
 * (29 arguments).
 *
 *   q    [total_q, num_heads,   head_size]   total_q = cu_seqlens_q[batch]
 *   k,v  [total_k, num_heads_k, head_size]   total_k = cu_seqlens_k[batch]
 *   cu_seqlens_q/k  device int32 [batch+1], exclusive prefix sums.
 *   softmax_lse_ptr fp32 [batch, num_heads, max_seqlen_q]; entries past a
 *        sequence's length are not written.
 *   mask_dims is `const void *` here, as in the reference (:66).
 */
bool flash_attn_varlen_fwd(const void * const q,
                           const void * const k,
                           const void * const v,
                           const int32_t * const cu_seqlens_q,
                           const int32_t * const cu_seqlens_k,
                           void * const rng_state,
                           void * const out,
                           void * const softmax_ptr,
                           void * const softmax_lse_ptr,
                           const int batch_size,
                           const int max_seqlen_q,
                           const int max_seqlen_k,
                           const int seqlen_q_rounded,
                           const int seqlen_k_rounded,
                           const int num_heads,
                           const int num_heads_k,
                           const int head_size,
                           const int head_size_rounded,
                           const float p_dropout,
                           const float softmax_scale,
                           const float softmax_unscale,
                           const bool is_causal,
                           const bool return_softmax,
                           const bool is_bf16,
                           fa_stream_t stream,
                           uint64_t seed,
                           uint64_t offset,
                           const void * const attn_mask,
                           const void * const mask_dims);

/*
 * Dense-batch backward. Replaces reference flash_attn.h:68-99 (32 arguments).
 *
 *   dout, q, out, dq   [batch, seqlen_q, num_heads,   head_size]
 *   k, v, dk, dv       [batch, seqlen_k, num_heads_k, head_size]
 *   softmax_lse  fp32 [batch, num_heads, seqlen_q] from the forward.
 *   softmax_d    fp32 scratch, >= batch*num_heads*seqlen_q_rounded elements;
 *                the library writes rowsum(dout*out) into it (const in the
 *                reference signature notwithstanding).
 *   dq_accum     fp32 scratch, >= batch*seqlen_q_rounded*num_heads*
 *                head_size_rounded elements.
 *   rng_state    unused by the backward (seed/offset come by value).
 *   num_splits   1 = deterministic dq accumulation; 0 = library's choice
 *                (atomics allowed); other values are treated as 0.
 *   seed/offset  must equal the forward's for the dropout mask to match.
 */
bool flash_attn_bwd(const void * const dout,
                    const void * const q,
                    const void * const k,
                    const void * const v,
                    const void * const out,
                    const void * const softmax_d,
                    const void * const softmax_lse,
                    void * const rng_state,
                    void * const dq,
                    void * const dk,
                    void * const dv,
                    void * const dq_accum,
                    const int batch_size,
                    const int seqlen_q,
                    const int seqlen_k,
                    const int seqlen_q_rounded,
                    const int seqlen_k_rounded,
                    const int num_heads,
                    const int num_heads_k,
                    const int head_size,
                    const int head_size_rounded,
                    const float p_dropout,
                    const float softmax_scale,
                    const float softmax_unscale,
                    const bool is_causal,
                    const bool is_bf16,
                    const int num_splits,
                    fa_stream_t stream,
                    uint64_t seed,
                    uint64_t offset,
                    const void * const attn_mask,
                    const int64_t * const mask_dims);

/*
 * Packed variable-length backward. Replaces reference flash_attn.h:101-134
 * (34 arguments). Layouts as flash_attn_varlen_fwd; scratch as flash_attn_bwd
 * with softmax_d / softmax_lse laid out [batch, num_heads, max_seqlen_q(_rounded)].
 */
bool flash_attn_varlen_bwd(const void * const dout,
                           const void * const q,
                           const void * const k,
                           const void * const v,
                           const void * const out,
                           const void * const softmax_d,
                           const void * const softmax_lse,
                           const int32_t * const cu_seqlens_q,
                           const int32_t * const cu_seqlens_k,
                           void * const rng_state,
                           void * const dq,
                           void * const dk,
                           void * const dv,
                           void * const dq_accum,
                           const int batch_size,
                           const int max_seqlen_q,
                           const int max_seqlen_k,
                           const int seqlen_q_rounded,
                           const int seqlen_k_rounded,
                           const int num_heads,
                           const int num_heads_k,
                           const int head_size,
                           const int head_size_rounded,
                           const float p_dropout,
                           const float softmax_scale,
                           const float softmax_unscale,
                           const bool is_causal,
                           const bool is_bf16,
                           const int num_splits,
                           fa_stream_t stream,
                           uint64_t seed,
                           uint64_t offset,
                           const void * attn_mask,
                           const int64_t * const mask_dims);

/* Stores a thread-local message (reference flash_attn.h:136). NULL clears it. */
void flash_attn_set_error(const char *msg);

/* Last message stored on this thread, never NULL (reference flash_attn.h:138). */
const char *flash_attn_error(void);

#ifdef __cplusplus
}
#endif

#endif /* FLASH_ATTN_B200_FLASH_ATTN_H_ */
