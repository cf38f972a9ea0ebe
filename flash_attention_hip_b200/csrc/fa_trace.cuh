// fa_trace.cuh -- optional in-kernel event timeline (build with -DFA_TRACE).
// One CTA (blockIdx == kTraceBlock) records (clock64 << 8 | event id) per warp
// role into a static device buffer; flash_attn_b200_trace_read() copies a
// role's events out.  Compiled out entirely in the product build.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace fa {
namespace trace {

constexpr int kRoles = 8;
constexpr int kMaxEvents = 4096;

#ifdef FA_TRACE
// internal linkage: each traced .cu file owns its buffer and its reader
static __device__ unsigned long long g_events[kRoles][kMaxEvents];
static __device__ int g_count[kRoles];

// Host: copies role `role`'s events to dst (synchronises the device); returns the count.
static inline int read_events(unsigned long long *dst, int role, int max_events) {
  int counts[kRoles];
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(counts, g_count, sizeof(counts)) != cudaSuccess) return -1;
  int n = counts[role] < max_events ? counts[role] : max_events;
  if (n > 0 && cudaMemcpyFromSymbol(dst, g_events, n * sizeof(unsigned long long),
                                    (size_t)role * kMaxEvents * sizeof(unsigned long long)) != cudaSuccess)
    return -1;
  return n;
}

struct Recorder {
  int role, n;
  bool on;
  __device__ __forceinline__ Recorder(int role_, bool on_) : role(role_), n(0), on(on_) { wall(254); }
  // (globaltimer ns << 8 | id): ids 254/255 bracket the recording so cycles convert to a clock rate
  __device__ __forceinline__ void wall(int id) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    if (on && n < kMaxEvents) g_events[role][n++] = (t << 8) | (unsigned)id;
    ev(id - 2);
  }
  __device__ __forceinline__ void ev(int id) {
    if (on && n < kMaxEvents) g_events[role][n++] = ((unsigned long long)clock64() << 8) | (unsigned)id;
  }
  __device__ __forceinline__ ~Recorder() {
    wall(255);
    if (on) g_count[role] = n;
  }
};
#define FA_TRACE_DECL(name, role, on) ::fa::trace::Recorder name((role), (on))
#define FA_TRACE_EV(name, id) name.ev(id)
#else
#define FA_TRACE_DECL(name, role, on)
#define FA_TRACE_EV(name, id)
#endif

}  // namespace trace
}  // namespace fa
