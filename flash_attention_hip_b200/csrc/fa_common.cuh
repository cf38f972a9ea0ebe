// fa_common.cuh -- parameter block and small device helpers shared by every
// kernel behind the flash_attn.h C ABI (include/flash_attn.h).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "philox_spec.h"

namespace fa {

// One attention problem, as the four ABI entry points describe it
// (reference flash_attn.h:10-36, :38-66, :68-99, :101-134).
struct Params {
  // forward tensors
  const void *q, *k, *v;
  void *out;
  float *lse;          // [b, h, seqlen_q]
  void *softmax_ret;   // [b, h, sq_rounded, sk_rounded] or null
  // backward tensors
  const void *dout;
  void *dq, *dk, *dv;
  float *softmax_d;    // [b, h, sq_rounded]
  float *dq_accum;     // [b, sq_rounded, h, d_rounded] scratch
  // packed layout (null = dense)
  const int32_t *cu_q, *cu_k;
  // optional additive bias
  const void *mask;
  int64_t mask_b, mask_h, mask_sq, mask_sk;

  int b, sq, sk;       // sq/sk: dense lengths or packed maxima
  int sq_rounded, sk_rounded;
  int h, hk, d, d_rounded;
  float scale;         // softmax_scale
  float p_dropout;
  float rp_dropout;    // 1/(1-p)
  uint32_t drop_threshold;  // keep iff byte < threshold
  uint64_t seed, offset;
  int batch_base;      // global index of local batch 0 (Philox head word; flash_attn_b200_set_batch_base)
  bool causal, is_bf16, deterministic;
};

// Per-sequence extents; dense and packed layouts differ only here.
struct SeqInfo {
  int q0, k0;   // first token row of this sequence in the [tokens, heads, d] view
  int sq, sk;   // its lengths
};

__device__ __forceinline__ SeqInfo seq_info(const Params &p, int b) {
  SeqInfo s;
  if (p.cu_q) {
    s.q0 = p.cu_q[b];
    s.k0 = p.cu_k[b];
    s.sq = p.cu_q[b + 1] - s.q0;
    s.sk = p.cu_k[b + 1] - s.k0;
  } else {
    s.q0 = b * p.sq;
    s.k0 = b * p.sk;
    s.sq = p.sq;
    s.sk = p.sk;
  }
  return s;
}

template <typename T> __device__ __forceinline__ float to_f32(T x);
template <> __device__ __forceinline__ float to_f32<__half>(__half x) { return __half2float(x); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }

template <typename T> __device__ __forceinline__ T from_f32(float x);
template <> __device__ __forceinline__ __half from_f32<__half>(float x) { return __float2half_rn(x); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- per-device launch state ------------------------------------------------------------------------
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and the SM count belong to a DEVICE, not to the process:
// a caller may drive several GPUs from one process (one host thread per device, cudaSetDevice before the
// call -- INTEGRATION.md), so both are cached per device ordinal, lock-free (re-doing either is harmless).
constexpr int kMaxCachedDevices = 64;

// SMs the persistent kernels leave free (thread-local; 0 outside the ring entry points).  A persistent launch holds one
// CTA on every SM until it is over, so a communication kernel that becomes runnable at the same moment finds no SM and
// runs AFTER it: the ring steps leave NCCL's send / receive CTAs a few SMs instead (csrc/fa_ring.cu).
inline int &persistent_sm_reserve() {
  static thread_local int reserve = 0;
  return reserve;
}

// Current device and the SM count the persistent kernels may fill.  Returns nullptr or a static error string.
inline const char *current_device(int *dev, int *sm_count) {
  static std::atomic<int> cache[kMaxCachedDevices];     // zero-initialised
  if (cudaGetDevice(dev) != cudaSuccess) return "cudaGetDevice failed";
  int c = (*dev >= 0 && *dev < kMaxCachedDevices) ? cache[*dev].load(std::memory_order_relaxed) : 0;
  if (c <= 0) {
    if (cudaDeviceGetAttribute(&c, cudaDevAttrMultiProcessorCount, *dev) != cudaSuccess || c <= 0)
      return "cudaDeviceGetAttribute(MultiProcessorCount) failed";
    if (*dev >= 0 && *dev < kMaxCachedDevices) cache[*dev].store(c, std::memory_order_relaxed);
  }
  const int r = persistent_sm_reserve();
  *sm_count = (r > 0 && c - r >= 16) ? c - r : c;
  return nullptr;
}

// Opt one kernel instantiation into `bytes` of dynamic shared memory on device `dev`, once per device.
// `done` is that instantiation's bit set (a function-local static of the launcher template).
template <typename Kernel>
inline const char *ensure_dynamic_smem(Kernel kern, int bytes, int dev, std::atomic<uint64_t> &done) {
  const bool cached = dev >= 0 && dev < kMaxCachedDevices;
  if (cached && ((done.load(std::memory_order_acquire) >> dev) & 1ull)) return nullptr;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
    return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed";
  if (cached) done.fetch_or(1ull << dev, std::memory_order_release);
  return nullptr;
}

// Work-item counters of the persistent kernels: a pool of self-resetting {next item, CTAs done} pairs in
// device memory (one pool per kernel family and device).  A launch takes the next slot round-robin; the
// last CTA to leave the item loop zeroes the pair again, so a launch needs no allocation, memset or free
// (the ABI has no argument for scratch of this kind).  Two launches share a slot only if kSchedSlots other
// launches of the same family were issued between them AND both are resident at once, which kernels that
// fill every SM cannot be.
constexpr int kSchedSlots = 64;
struct SchedSlot { int next, done, pad[2]; };
inline int next_sched_slot(std::atomic<unsigned> &ticket) {
  return (int)(ticket.fetch_add(1u, std::memory_order_relaxed) % (unsigned)kSchedSlots);
}
// Host: this launch's slot of the pool behind device symbol `pool_symbol` (address cached per device).
template <typename Symbol>
inline const char *take_sched_slot(const Symbol &pool_symbol, int dev, std::atomic<SchedSlot *> *cache,
                                   std::atomic<unsigned> &ticket, SchedSlot **out) {
  const bool cached = dev >= 0 && dev < kMaxCachedDevices;
  SchedSlot *pool = cached ? cache[dev].load(std::memory_order_acquire) : nullptr;
  if (!pool) {
    if (cudaGetSymbolAddress(reinterpret_cast<void **>(&pool), pool_symbol) != cudaSuccess || !pool)
      return "cudaGetSymbolAddress(item counters) failed";
    if (cached) cache[dev].store(pool, std::memory_order_release);
  }
  *out = pool + next_sched_slot(ticket);
  return nullptr;
}
#ifdef __CUDACC__
// Scheduler lane of a persistent CTA: the next item, or -1 once the items have run out (called exactly
// once with that outcome per CTA: the last CTA to see it re-arms the slot for a later launch).
__device__ __forceinline__ int sched_next_item(SchedSlot *slot, int n_items) {
  const int L = atomicAdd(&slot->next, 1);
  if (L < n_items) return L;
  __threadfence();                                  // my last pull is ordered before my `done` count
  if (atomicAdd(&slot->done, 1) == (int)gridDim.x - 1) {
    slot->next = 0;
    slot->done = 0;
    __threadfence();
  }
  return -1;
}
#endif

// Launchers (defined in the .cu files; all asynchronous on `stream`).
const char *launch_generic_fwd(const Params &p, cudaStream_t stream);
const char *launch_generic_bwd(const Params &p, cudaStream_t stream);
// softmax_d = rowsum(dout o out); also zeroes `zero_floats` floats at `zero_acc` (the tcgen05 backward's dq_accum) when non-null
const char *launch_bwd_preprocess(const Params &p, cudaStream_t stream, float *zero_acc, long long zero_floats);
// floats of dq_accum the tcgen05 backward uses: [b][h][query tiles][ceil(d/32)][128][32]
long long tc_bwd_scratch_floats(const Params &p);
bool tc_bwd_uses_scratch(const Params &p);
// all-shapes backward on warp-level MMAs (fa_mma_bwd.cu); what launch_generic_bwd runs
const char *launch_mma_bwd(const Params &p, cudaStream_t stream);

// tcgen05 paths; return nullptr on success, else a static error string.
bool tc_fwd_supported(const Params &p);
const char *launch_tc_fwd(const Params &p, cudaStream_t stream);     // ignores p.softmax_ret
// S_dmask (p.softmax_ret) from q, k, the bias and the LSE a forward has written (fa_softmax_ret.cu)
const char *launch_softmax_ret(const Params &p, cudaStream_t stream);
bool tc_bwd_supported(const Params &p);
const char *launch_tc_bwd(const Params &p, cudaStream_t stream);
const char *launch_tc_bwd_wide(const Params &p, cudaStream_t stream);   // head sizes 136..256 (fa_tc_bwd_wide.cu)

// Ring context-parallel combine kernels (fa_cp.cu).
const char *launch_cp_merge(void *o_acc, void *lse_acc, const void *o_blk, const void *lse_blk, int rows, int h,
                            int d, int64_t lse_acc_stride, int64_t lse_blk_stride, bool is_bf16,
                            cudaStream_t stream);
const char *launch_cp_accumulate(void *acc, const void *blk, int64_t n, bool is_bf16, cudaStream_t stream);
const char *launch_cp_store(void *dst, const void *src, int64_t n, bool is_bf16, cudaStream_t stream);

// Event timelines of a -DFA_TRACE build (0 events otherwise); see fa_trace.cuh.
int trace_read_bwd(unsigned long long *dst, int role, int max_events);
int trace_read_fwd(unsigned long long *dst, int role, int max_events);

}  // namespace fa
