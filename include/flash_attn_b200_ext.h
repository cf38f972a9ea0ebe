/*
 * flash_attn_b200_ext.h -- additions beside the six reference symbols.
 *
 * Nothing here is needed by a caller of the reference ABI (include/flash_attn.h);
 * the reference header (/root/reference/flash_attn.h) declares none of it.
 * These entry points exist for tests, benchmarks and the multi-GPU drivers.
 */
#ifndef FLASH_ATTN_B200_EXT_H_
#define FLASH_ATTN_B200_EXT_H_

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#include "flash_attn.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Kernel families behind the ABI. */
enum {
  FA_B200_PATH_AUTO = 0,    /* tcgen05 where it applies, else the all-shapes kernels */
  FA_B200_PATH_GENERIC = 1, /* force the all-shapes SIMT kernels */
  FA_B200_PATH_TCGEN05 = 2  /* require the tcgen05 kernels; fail if they do not apply */
};

/* Thread-local kernel-family selection for the next calls (default AUTO). */
void flash_attn_b200_set_path(int path);

/* Which family the last successful call on this thread launched
 * (FA_B200_PATH_GENERIC or FA_B200_PATH_TCGEN05; 0 before any call). */
int flash_attn_b200_last_path(void);

/* Batch sharding.  The dropout stream is indexed by (GLOBAL batch, head, row, column); a rank that
 * holds batches [base, base + n) of a global batch passes `batch_base = base` so that its masks equal the
 * unsharded call's.  The four functions below are the reference entry points (include/flash_attn.h, reference
 * flash_attn.h:10-134) with that ONE extra leading argument and otherwise identical arguments, results and
 * error behaviour; batch_base = 0 is exactly the reference call.  A forward and its backward must be given the
 * same batch_base.  (Round 1 had a sticky thread-local setter instead; it is gone: hidden per-thread state
 * beside the ABI is a trap for callers that run on thread pools.) */
bool flash_attn_b200_fwd_shard(int batch_base, const void *q, const void *k, const void *v, void *rng_state, void *out,
                               void *softmax_ptr, void *softmax_lse_ptr, int batch_size, int seqlen_q, int seqlen_k,
                               int seqlen_q_rounded, int seqlen_k_rounded, int num_heads, int num_heads_k, int head_size,
                               int head_size_rounded, float p_dropout, float softmax_scale, float softmax_unscale,
                               bool is_causal, bool return_softmax, bool is_bf16, fa_stream_t stream, uint64_t seed,
                               uint64_t offset, const void *attn_mask, const int64_t *mask_dims);
bool flash_attn_b200_varlen_fwd_shard(int batch_base, const void *q, const void *k, const void *v,
                                      const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, void *rng_state,
                                      void *out, void *softmax_ptr, void *softmax_lse_ptr, int batch_size,
                                      int max_seqlen_q, int max_seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded,
                                      int num_heads, int num_heads_k, int head_size, int head_size_rounded,
                                      float p_dropout, float softmax_scale, float softmax_unscale, bool is_causal,
                                      bool return_softmax, bool is_bf16, fa_stream_t stream, uint64_t seed,
                                      uint64_t offset, const void *attn_mask, const void *mask_dims);
bool flash_attn_b200_bwd_shard(int batch_base, const void *dout, const void *q, const void *k, const void *v,
                               const void *out, const void *softmax_d, const void *softmax_lse, void *rng_state,
                               void *dq, void *dk, void *dv, void *dq_accum, int batch_size, int seqlen_q, int seqlen_k,
                               int seqlen_q_rounded, int seqlen_k_rounded, int num_heads, int num_heads_k, int head_size,
                               int head_size_rounded, float p_dropout, float softmax_scale, float softmax_unscale,
                               bool is_causal, bool is_bf16, int num_splits, fa_stream_t stream, uint64_t seed,
                               uint64_t offset, const void *attn_mask, const int64_t *mask_dims);
bool flash_attn_b200_varlen_bwd_shard(int batch_base, const void *dout, const void *q, const void *k, const void *v,
                                      const void *out, const void *softmax_d, const void *softmax_lse,
                                      const int32_t *cu_seqlens_q, const int32_t *cu_seqlens_k, void *rng_state,
                                      void *dq, void *dk, void *dv, void *dq_accum, int batch_size, int max_seqlen_q,
                                      int max_seqlen_k, int seqlen_q_rounded, int seqlen_k_rounded, int num_heads,
                                      int num_heads_k, int head_size, int head_size_rounded, float p_dropout,
                                      float softmax_scale, float softmax_unscale, bool is_causal, bool is_bf16,
                                      int num_splits, fa_stream_t stream, uint64_t seed, uint64_t offset,
                                      const void *attn_mask, const int64_t *mask_dims);

/* Dense backward for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference header.
 * Same results as flash_attn_bwd (include/flash_attn.h) with num_splits = 0 and no attn_mask, but the two scratch
 * tensors (softmax_d, dq_accum) are not arguments: they live in one static device array of
 * FA_B200_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit.  The call performs no
 * cudaMalloc / cudaFree / stream-ordered allocation.  Limit: 4 * batch * heads * seqlen_q_rounded *
 * (1 + head_size_rounded) bytes (seqlen_q rounded up to 128, head_size to 32; e.g. B2 H8 S1024 D128 = 8.1 MiB)
 * must not exceed FA_B200_STATIC_SCRATCH_BYTES, else the call returns false with a message and launches nothing.
 * Every call on a device shares the array: what must start at zero is cleared by the call itself, and two calls
 * that overlap on the GPU (different streams, or two host threads on one device) race on it -- serialise them. */
#define FA_B200_STATIC_SCRATCH_BYTES (32u << 20)
bool flash_attn_b200_bwd_static_scratch(const void *dout, const void *q, const void *k, const void *v,
                                        const void *out, const void *softmax_lse, void *dq, void *dk, void *dv,
                                        int batch_size, int seqlen_q, int seqlen_k, int num_heads, int num_heads_k,
                                        int head_size, float p_dropout, float softmax_scale, bool is_causal,
                                        bool is_bf16, fa_stream_t stream, uint64_t seed, uint64_t offset);

/* Number of kernel launches the last successful call on this thread enqueued. */
int flash_attn_b200_last_launch_count(void);

/* ---- ring context parallel (BASELINE.json config 5) -------------------------------------------
 * Every ring step is an ordinary flash_attn_fwd / flash_attn_bwd call on a (query slice, visiting
 * K/V chunk) pair; these HBM-bound helpers combine the per-step results.  All pointers are device
 * memory, all work is enqueued on `stream`; false + flash_attn_error() on failure.
 *
 * cp_merge: (o_acc fp32 [rows, h, d], lse_acc fp32, element (head, row) at head*lse_acc_stride+row)
 *   <- log-sum-exp merge with one step's (o_blk 16-bit [rows, h, d], lse_blk fp32, stride
 *   lse_blk_stride).  lse == +inf marks a row that has seen no key yet (initialise lse_acc so). */
bool flash_attn_b200_cp_merge(void *o_acc, void *lse_acc, const void *o_blk, const void *lse_blk, int rows,
                              int num_heads, int head_size, int64_t lse_acc_stride, int64_t lse_blk_stride,
                              bool is_bf16, fa_stream_t stream);
/* acc_f32[i] += blk[i] (16-bit), i < n: dq / dk / dv partial sums of the ring backward. */
bool flash_attn_b200_cp_accumulate(void *acc_f32, const void *blk, int64_t n, bool is_bf16, fa_stream_t stream);
/* dst[i] (16-bit) = src_f32[i], i < n. */
bool flash_attn_b200_cp_store(void *dst, const void *src_f32, int64_t n, bool is_bf16, fa_stream_t stream);

/* ---- ring context parallel as a library call ------------------------------------------------------
 * No counterpart in the reference header.  One process (or host thread) per GPU; the N ranks of one NVLink box each
 * hold a shard of the sequence: q [b, s_local, h, d], k / v [b, s_local, hk, d] (causal: the ZIG-ZAG shard -- chunks
 * r and 2N-1-r of the 2N equal chunks of the full sequence, s_local even; non-causal: any contiguous split).
 * ring_fwd returns this rank's rows of full-sequence attention: out [b, s_local, h, d], lse fp32 [b, h, s_local].
 * ring_bwd takes the forward's (out, lse) and returns dq, dk, dv of the shard.  K/V chunks (and, in the backward,
 * the fp32 dK/dV partial sums) travel round the ring with ncclSend / ncclRecv on a communication stream owned by the
 * comm object, one step ahead of the kernels that consume them; everything else is enqueued on `stream`.  Every
 * rank must make the same call.  No attn_mask in this mode.  NCCL is bound at run time (dlopen of
 * libnccl.so.2): nothing else in this library needs it.
 *
 * Dropout (the *_dropout pair; p_dropout = 0 is exactly ring_fwd / ring_bwd): the step of rank r on the K/V chunk of
 * rank src is an ordinary call with the caller's seed, offset + ((r * N + src) << 48) and batch_base = the batch entry,
 * rows and columns counted inside that step's (query slice, K/V chunk) -- see csrc/fa_ring.cu for the slices.  The mask
 * is therefore a function of (seed, offset, N) and NOT the unsharded call's (as in every ring attention built on
 * per-block calls); forward and backward of one pair draw the same mask; a checker can replay any rank's steps with the
 * ordinary entry points on one GPU (tests/test_gpu_shard_ring.py).
 *
 * Bootstrap like NCCL's own: one rank calls ring_unique_id, hands the FA_B200_RING_ID_BYTES to the others by any
 * means, every rank (its device current) calls ring_comm_create.  ring_comm_from_nccl borrows a communicator the
 * caller already has (not destroyed by ring_comm_destroy).
 * `workspace`: device memory of at least ring_*_workspace_bytes(...), 256-byte aligned, owned by the caller. */
#define FA_B200_RING_ID_BYTES 128
typedef struct fa_ring_comm fa_ring_comm_t;
bool flash_attn_b200_ring_unique_id(void *id128);
bool flash_attn_b200_ring_comm_create(const void *id128, int nranks, int rank, fa_ring_comm_t **comm);
bool flash_attn_b200_ring_comm_from_nccl(void *nccl_comm, int nranks, int rank, fa_ring_comm_t **comm);
void flash_attn_b200_ring_comm_destroy(fa_ring_comm_t *comm);
size_t flash_attn_b200_ring_fwd_workspace_bytes(int nranks, int b, int s_local, int h, int hk, int d);
size_t flash_attn_b200_ring_bwd_workspace_bytes(int nranks, int b, int s_local, int h, int hk, int d);
bool flash_attn_b200_ring_fwd(fa_ring_comm_t *comm, const void *q, const void *k, const void *v, void *out, void *lse,
                              void *workspace, size_t workspace_bytes, int b, int s_local, int h, int hk, int d,
                              float softmax_scale, bool is_causal, bool is_bf16, fa_stream_t stream);
bool flash_attn_b200_ring_bwd(fa_ring_comm_t *comm, const void *dout, const void *q, const void *k, const void *v,
                              const void *out, const void *lse, void *dq, void *dk, void *dv, void *workspace,
                              size_t workspace_bytes, int b, int s_local, int h, int hk, int d, float softmax_scale,
                              bool is_causal, bool is_bf16, fa_stream_t stream);
bool flash_attn_b200_ring_fwd_dropout(fa_ring_comm_t *comm, const void *q, const void *k, const void *v, void *out,
                                      void *lse, void *workspace, size_t workspace_bytes, int b, int s_local, int h, int hk,
                                      int d, float softmax_scale, bool is_causal, bool is_bf16, fa_stream_t stream,
                                      float p_dropout, uint64_t seed, uint64_t offset);
bool flash_attn_b200_ring_bwd_dropout(fa_ring_comm_t *comm, const void *dout, const void *q, const void *k, const void *v,
                                      const void *out, const void *lse, void *dq, void *dk, void *dv, void *workspace,
                                      size_t workspace_bytes, int b, int s_local, int h, int hk, int d, float softmax_scale,
                                      bool is_causal, bool is_bf16, fa_stream_t stream, float p_dropout, uint64_t seed,
                                      uint64_t offset);

/* Developer builds only (-DFA_TRACE): copies the in-kernel event timeline of warp role `role`
 * of the forward (kernel 0) or backward (kernel 1) tcgen05 kernel's traced CTA into dst as
 * (clock64 << 8 | event id) words; synchronises the device.  Product builds return 0. */
int flash_attn_b200_trace_read(int kernel, int role, unsigned long long *dst, int max_events);

/* Library version string, e.g. "flash_attn_b200 0.1 (sm_100a)". */
const char *flash_attn_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif
