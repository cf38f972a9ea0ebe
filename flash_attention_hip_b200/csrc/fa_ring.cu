// fa_ring.cu -- ring context parallelism as C entry points (include/flash_attn_b200_ext.h): BASELINE.json
// config 5, "long-context S=128k D=128 causal ring context-parallel over NVLink".  The one place in the
// library with an exchange step.  No counterpart in the reference header (flash_attn.h declares four
// single-GPU calls); every ring step IS one of those calls -- flash_attn_fwd / flash_attn_bwd
// (reference flash_attn.h:10-36, :68-99) on a (query slice, visiting K/V chunk) pair -- and this file is
// the host schedule around them:
//   * the sequence is split over the N ranks of one NVLink box; K/V chunks travel round the ring with
//     ncclSend / ncclRecv on a communication stream of the ring's own, ONE STEP AHEAD of the kernels that
//     consume them (event-chained to the caller's stream), so the transfer is hidden under the tile loop;
//   * causal attention uses zig-zag sharding (rank r holds chunks r and 2N-1-r of 2N), which makes every
//     step one of three dense calls (c = chunk rows, src = owner of the visiting K/V):
//       src == r : causal,     q[0:2c] x kv[0:2c]
//       src <  r : non-causal, q[0:2c] x kv[0:c]
//       src >  r : non-causal, q[c:2c] x kv[0:2c]
//   * forward: each step's (out, lse) is kept (16-bit out, fp32 lse) and ONE kernel merges the N partial
//     results at the end (log-sum-exp weights) -- 2 bytes read per element and step instead of the 10 an
//     fp32 running accumulator costs;
//   * backward: dK / dV partial sums travel with their K/V chunk one hop behind it, in fp32 (N - 1 roundings
//     to 16 bits on the way would cost accuracy; the 2 x 67 MB per step at S = 128k, N = 8 are ~0.15 ms of
//     NVLink time under ~6 ms of kernels), and arrive home after the last step.
// NCCL is bound at run time (dlopen of libnccl.so.2 -- the one already in the process when the caller is a
// torch program): callers of the reference ABI never need it.  Dropout: the *_dropout entry points (step_offset below).
#include <dlfcn.h>
#include <nccl.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>

#include "../../include/flash_attn.h"
#include "../../include/flash_attn_b200_ext.h"
#include "fa_common.cuh"

namespace {

constexpr int kMaxRanks = 16;

struct Nccl {
  decltype(&ncclGetUniqueId) GetUniqueId;
  decltype(&ncclCommInitRank) CommInitRank;
  decltype(&ncclCommDestroy) CommDestroy;
  decltype(&ncclSend) Send;
  decltype(&ncclRecv) Recv;
  decltype(&ncclGroupStart) GroupStart;
  decltype(&ncclGroupEnd) GroupEnd;
  decltype(&ncclGetErrorString) GetErrorString;
  bool ok = false;
};

const Nccl &nccl() {
  static Nccl n;
  static std::once_flag once;
  std::call_once(once, [] {
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return;
#define FA_SYM(name) n.name = reinterpret_cast<decltype(n.name)>(dlsym(h, "nccl" #name)); if (!n.name) return
    FA_SYM(GetUniqueId); FA_SYM(CommInitRank); FA_SYM(CommDestroy); FA_SYM(Send); FA_SYM(Recv);
    FA_SYM(GroupStart); FA_SYM(GroupEnd); FA_SYM(GetErrorString);
#undef FA_SYM
    n.ok = true;
  });
  return n;
}

bool failf(const char *fmt, ...) __attribute__((format(printf, 1, 2)));
bool failf(const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  flash_attn_set_error(buf);
  return false;
}

#define FA_CUDA(call)                                                                      \
  do {                                                                                     \
    const cudaError_t e_ = (call);                                                         \
    if (e_ != cudaSuccess) return failf("%s: %s: %s", fn, #call, cudaGetErrorString(e_));  \
  } while (0)
#define FA_NCCL(call)                                                                      \
  do {                                                                                     \
    const ncclResult_t r_ = (call);                                                        \
    if (r_ != ncclSuccess) return failf("%s: %s: %s", fn, #call, nccl().GetErrorString(r_)); \
  } while (0)
#define FA_CALL(call)                                                                      \
  do {                                                                                     \
    if (!(call)) return false; /* the callee has set the message */                        \
  } while (0)

size_t up256(size_t x) { return (x + 255) / 256 * 256; }

// One ring step: (q rows, kv rows, causal flag) in local rows.
struct Step { int q0, q1, k0, k1; bool causal; };
Step ring_step(bool causal, int rank, int src, int s_local) {
  if (!causal) return {0, s_local, 0, s_local, false};
  const int c = s_local / 2;
  if (src == rank) return {0, s_local, 0, s_local, true};
  if (src < rank) return {0, s_local, 0, c, false};
  return {c, s_local, 0, s_local, false};
}

// Dropout in ring mode (include/flash_attn_b200_ext.h): every step is an ORDINARY call with the caller's seed, its own
// Philox offset and batch_base = the batch entry -- rows and columns counted inside the step's (query slice, K/V chunk).
// The pair index goes into bits 48.. of the offset, i.e. into the Philox KEY (csrc/philox_spec.h: key0 = lo(seed) ^
// hi(offset)): the N * N pairs draw from unrelated streams, the backward of a pair repeats its forward's.
uint64_t step_offset(uint64_t offset, int rank, int src, int nranks) {
  return offset + ((uint64_t)(rank * nranks + src) << 48);
}

// ---- final merge of the forward: N partial (out, lse) -> out, lse ---------------------------------------
struct MergeArgs {
  const void *o[kMaxRanks];       // step t: [b][rows_t][h][d], 16-bit
  const float *lse[kMaxRanks];    // step t: [b][h][rows_t]
  int q0[kMaxRanks], rows[kMaxRanks];
  int n;
};

// One warp per (batch, row, head).  Lane t holds step t's LSE for the row (+inf: the step did not cover it, or saw no key):
// warp max / sum give the log-sum-exp weights; every lane then owns 4 output elements per 128 (8-byte loads and stores).
template <typename T>
__global__ void __launch_bounds__(256)
ring_merge_kernel(const MergeArgs m, T *__restrict__ out, float *__restrict__ lse_out, int b, int s, int h, int d) {
  const unsigned warp = (unsigned)((blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5);   // b * s * h < 2^32 (checked by the caller)
  const int lane = threadIdx.x & 31;
  if (warp >= (unsigned)b * (unsigned)s * (unsigned)h) return;
  const int head = (int)(warp % (unsigned)h), row = (int)((warp / (unsigned)h) % (unsigned)s), bi = (int)(warp / ((unsigned)h * (unsigned)s));
  float l = INFINITY;
  int rel = 0;
  if (lane < m.n) {
    rel = row - m.q0[lane];
    if (rel >= 0 && rel < m.rows[lane]) l = m.lse[lane][((size_t)bi * h + head) * m.rows[lane] + rel];
  }
  const float mx = fa::warp_max(l == INFINITY ? -INFINITY : l);
  T *dst = out + (((size_t)bi * s + row) * h + head) * d;
  if (mx == -INFINITY) {                                 // no key anywhere: out = 0, lse = +inf
    for (int c = lane * 4; c < d; c += 128) *reinterpret_cast<uint2 *>(dst + c) = make_uint2(0u, 0u);
    if (lane == 0) lse_out[((size_t)bi * h + head) * s + row] = INFINITY;
    return;
  }
  const float w = l == INFINITY ? 0.f : __expf(l - mx);  // my step's weight, unnormalised
  const float tot = fa::warp_sum(w);
  const float inv = 1.f / tot;
  for (int c = lane * 4; c < d; c += 128) {              // d is a multiple of 8: a lane's 4 elements never straddle the end
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    for (int t = 0; t < m.n; ++t) {
      const float wt = __shfl_sync(0xffffffffu, w, t);
      const int rt = __shfl_sync(0xffffffffu, rel, t);
      if (wt != 0.f) {                                   // (warp-uniform)
        const T *src = static_cast<const T *>(m.o[t]) + (((size_t)bi * m.rows[t] + rt) * h + head) * d + c;
        const uint2 v = *reinterpret_cast<const uint2 *>(src);
        const T *e = reinterpret_cast<const T *>(&v);
        a0 = fmaf(wt, fa::to_f32<T>(e[0]), a0); a1 = fmaf(wt, fa::to_f32<T>(e[1]), a1);
        a2 = fmaf(wt, fa::to_f32<T>(e[2]), a2); a3 = fmaf(wt, fa::to_f32<T>(e[3]), a3);
      }
    }
    T o4[4] = {fa::from_f32<T>(a0 * inv), fa::from_f32<T>(a1 * inv), fa::from_f32<T>(a2 * inv), fa::from_f32<T>(a3 * inv)};
    *reinterpret_cast<uint2 *>(dst + c) = *reinterpret_cast<const uint2 *>(o4);
  }
  if (lane == 0) lse_out[((size_t)bi * h + head) * s + row] = mx + __logf(tot);
}

}  // namespace

struct fa_ring_comm {
  ncclComm_t comm = nullptr;
  bool owns_comm = false;
  int nranks = 1, rank = 0, device = 0;
  cudaStream_t cs = nullptr;                     // the ring's communication stream
  cudaEvent_t ev_in = nullptr;                   // the caller's stream has reached the call (inputs are ready)
  cudaEvent_t ev_xchg[kMaxRanks + 1] = {};       // K/V of step t has arrived
  cudaEvent_t ev_done[kMaxRanks + 1] = {};       // the kernels of step t have been enqueued ... and finished
  cudaEvent_t ev_acc[kMaxRanks + 1] = {};        // backward: step t's dK/dV have been added to the travelling sums
  cudaEvent_t ev_dkv[kMaxRanks + 1] = {};        // backward: the sums sent after step t have arrived
};

namespace {

bool make_events(const char *fn, fa_ring_comm *c) {
  FA_CUDA(cudaGetDevice(&c->device));
  // HIGHEST priority: when an exchange and the step's attention kernel become runnable at the same moment, the NCCL
  // kernel's few CTAs must be placed first -- the attention kernels are persistent (one CTA per SM until the launch is
  // over), so an exchange that loses that race runs AFTER the step instead of under it (measured at N = 8: 19.1 ms forward
  // at S = 128k with a default-priority stream).
  int prio_least = 0, prio_greatest = 0;
  FA_CUDA(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
  FA_CUDA(cudaStreamCreateWithPriority(&c->cs, cudaStreamNonBlocking, prio_greatest));
  FA_CUDA(cudaEventCreateWithFlags(&c->ev_in, cudaEventDisableTiming));
  for (int t = 0; t <= kMaxRanks; ++t) {
    FA_CUDA(cudaEventCreateWithFlags(&c->ev_xchg[t], cudaEventDisableTiming));
    FA_CUDA(cudaEventCreateWithFlags(&c->ev_done[t], cudaEventDisableTiming));
    FA_CUDA(cudaEventCreateWithFlags(&c->ev_acc[t], cudaEventDisableTiming));
    FA_CUDA(cudaEventCreateWithFlags(&c->ev_dkv[t], cudaEventDisableTiming));
  }
  return true;
}

// send `bytes` at `src` to the next rank and receive as many into `dst` from the previous one (comm stream)
bool exchange(const char *fn, fa_ring_comm *c, const void *src0, void *dst0, size_t bytes0, const void *src1, void *dst1,
              size_t bytes1) {
  const int nxt = (c->rank + 1) % c->nranks, prv = (c->rank + c->nranks - 1) % c->nranks;
  FA_NCCL(nccl().GroupStart());
  FA_NCCL(nccl().Send(src0, bytes0, ncclInt8, nxt, c->comm, c->cs));
  FA_NCCL(nccl().Recv(dst0, bytes0, ncclInt8, prv, c->comm, c->cs));
  if (bytes1) {
    FA_NCCL(nccl().Send(src1, bytes1, ncclInt8, nxt, c->comm, c->cs));
    FA_NCCL(nccl().Recv(dst1, bytes1, ncclInt8, prv, c->comm, c->cs));
  }
  FA_NCCL(nccl().GroupEnd());
  return true;
}

struct RingShape { int b, s, h, hk, d; bool bf16; };

// SMs left to NCCL's send / receive kernel while a ring step's (persistent) attention kernels run.  Without them the
// exchange of step t + 1, which becomes runnable at the same moment as the kernels of step t, finds every SM taken and
// runs after them: measured at N = 8, S = 128k: 18.8 ms forward with no SM left free (13.5 ms of kernels + 8 x 0.66 ms of
// exposed exchange).  FA_B200_RING_SM_RESERVE overrides the default.
int ring_sm_reserve(int nranks) {
  if (nranks <= 1) return 0;
  static const int v = [] {
    const char *e = getenv("FA_B200_RING_SM_RESERVE");
    return e ? atoi(e) : 8;
  }();
  return v < 0 ? 0 : v;
}
struct ReserveGuard {
  int saved;
  explicit ReserveGuard(int r) : saved(fa::persistent_sm_reserve()) { fa::persistent_sm_reserve() = r; }
  ~ReserveGuard() { fa::persistent_sm_reserve() = saved; }
};

bool check_shape(const char *fn, const fa_ring_comm *c, const RingShape &r, bool causal) {
  if (!c) return failf("%s: comm is NULL", fn);
  if (r.b <= 0 || r.s <= 0 || r.h <= 0 || r.hk <= 0 || r.h % r.hk) return failf("%s: bad shape", fn);
  if (r.d <= 0 || r.d > 256 || r.d % 8) return failf("%s: head_size must be a multiple of 8 in [8, 256]", fn);
  if (causal && r.s % 2) return failf("%s: causal ring attention needs an even local length (zig-zag chunks)", fn);
  if (c->nranks > kMaxRanks) return failf("%s: at most %d ranks", fn, kMaxRanks);
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev != c->device)
    return failf("%s: the communicator belongs to device %d, the current device is %d", fn, c->device, dev);
  return true;
}

}  // namespace

extern "C" {

bool flash_attn_b200_ring_unique_id(void *id128) {
  const char *fn = "flash_attn_b200_ring_unique_id";
  if (!id128) return failf("%s: NULL", fn);
  if (!nccl().ok) return failf("%s: libnccl.so.2 could not be loaded", fn);
  static_assert(sizeof(ncclUniqueId) == FA_B200_RING_ID_BYTES, "ncclUniqueId size");
  FA_NCCL(nccl().GetUniqueId(static_cast<ncclUniqueId *>(id128)));
  return true;
}

bool flash_attn_b200_ring_comm_create(const void *id128, int nranks, int rank, fa_ring_comm_t **out) {
  const char *fn = "flash_attn_b200_ring_comm_create";
  if (!out || nranks < 1 || rank < 0 || rank >= nranks || (nranks > 1 && !id128)) return failf("%s: bad argument", fn);
  if (nranks > kMaxRanks) return failf("%s: at most %d ranks", fn, kMaxRanks);
  fa_ring_comm *c = new (std::nothrow) fa_ring_comm;
  if (!c) return failf("%s: out of memory", fn);
  c->nranks = nranks; c->rank = rank;
  if (nranks > 1) {
    if (!nccl().ok) { delete c; return failf("%s: libnccl.so.2 could not be loaded", fn); }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    const ncclResult_t r = nccl().CommInitRank(&c->comm, nranks, id, rank);
    if (r != ncclSuccess) { delete c; return failf("%s: ncclCommInitRank: %s", fn, nccl().GetErrorString(r)); }
    c->owns_comm = true;
  }
  if (!make_events(fn, c)) { flash_attn_b200_ring_comm_destroy(c); return false; }
  *out = c;
  return true;
}

bool flash_attn_b200_ring_comm_from_nccl(void *nccl_comm, int nranks, int rank, fa_ring_comm_t **out) {
  const char *fn = "flash_attn_b200_ring_comm_from_nccl";
  if (!out || !nccl_comm || nranks < 1 || rank < 0 || rank >= nranks) return failf("%s: bad argument", fn);
  if (nranks > kMaxRanks) return failf("%s: at most %d ranks", fn, kMaxRanks);
  if (!nccl().ok) return failf("%s: libnccl.so.2 could not be loaded", fn);
  fa_ring_comm *c = new (std::nothrow) fa_ring_comm;
  if (!c) return failf("%s: out of memory", fn);
  c->comm = static_cast<ncclComm_t>(nccl_comm); c->nranks = nranks; c->rank = rank;
  if (!make_events(fn, c)) { flash_attn_b200_ring_comm_destroy(c); return false; }
  *out = c;
  return true;
}

void flash_attn_b200_ring_comm_destroy(fa_ring_comm_t *c) {
  if (!c) return;
  if (c->cs) { cudaStreamSynchronize(c->cs); }
  if (c->comm && c->owns_comm) nccl().CommDestroy(c->comm);
  if (c->ev_in) cudaEventDestroy(c->ev_in);
  for (int t = 0; t <= kMaxRanks; ++t) {
    if (c->ev_xchg[t]) cudaEventDestroy(c->ev_xchg[t]);
    if (c->ev_done[t]) cudaEventDestroy(c->ev_done[t]);
    if (c->ev_acc[t]) cudaEventDestroy(c->ev_acc[t]);
    if (c->ev_dkv[t]) cudaEventDestroy(c->ev_dkv[t]);
  }
  if (c->cs) cudaStreamDestroy(c->cs);
  delete c;
}

// Workspace layouts (all regions 256-byte aligned):
//   forward : R[2] visiting K|V, then per step t: out_t [b][s][h][d] 16-bit, lse_t [b][h][s] fp32
//   backward: R[2]; D[2] travelling dK|dV fp32; dq_acc fp32; dq_t, dk_t, dv_t 16-bit (one batch entry);
//             lse_c, softmax_d, dq_accum (one batch entry)
size_t flash_attn_b200_ring_fwd_workspace_bytes(int nranks, int b, int s_local, int h, int hk, int d) {
  const size_t kv = up256((size_t)2 * b * s_local * hk * d * 2);
  const size_t o = up256((size_t)b * s_local * h * d * 2), l = up256((size_t)b * h * s_local * 4);
  return (nranks > 1 ? 2 * kv : 0) + (size_t)nranks * (o + l);
}

size_t flash_attn_b200_ring_bwd_workspace_bytes(int nranks, int b, int s_local, int h, int hk, int d) {
  const size_t s_r = ((size_t)s_local + 127) / 128 * 128, d_r = ((size_t)d + 31) / 32 * 32;
  const size_t kv = up256((size_t)2 * b * s_local * hk * d * 2), dkv = up256((size_t)2 * b * s_local * hk * d * 4);
  return (nranks > 1 ? 2 * kv : 0) + (nranks > 1 ? 2 : 1) * dkv + up256((size_t)b * s_local * h * d * 4) +
         up256((size_t)s_local * h * d * 2) + 2 * up256((size_t)s_local * hk * d * 2) + up256((size_t)h * s_local * 4) +
         up256((size_t)h * s_r * 4) + up256(s_r * h * d_r * 4);
}

bool flash_attn_b200_ring_fwd_dropout(fa_ring_comm_t *c, const void *q, const void *k, const void *v, void *out, void *lse,
                                      void *workspace, size_t workspace_bytes, int b, int s_local, int h, int hk, int d,
                                      float softmax_scale, bool is_causal, bool is_bf16, fa_stream_t stream,
                                      float p_dropout, uint64_t seed, uint64_t offset) {
  const char *fn = "flash_attn_b200_ring_fwd";
  const RingShape r{b, s_local, h, hk, d, is_bf16};
  FA_CALL(check_shape(fn, c, r, is_causal));
  if (!(p_dropout >= 0.f && p_dropout <= 1.f)) return failf("%s: p_dropout must be in [0, 1]", fn);
  if (!q || !k || !v || !out || !lse || !workspace) return failf("%s: NULL pointer", fn);
  const int N = c->nranks;
  if (workspace_bytes < flash_attn_b200_ring_fwd_workspace_bytes(N, b, s_local, h, hk, d))
    return failf("%s: workspace too small (flash_attn_b200_ring_fwd_workspace_bytes)", fn);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t kv_elems = (size_t)b * s_local * hk * d;          // of K (and of V)
  const size_t kv_bytes = kv_elems * 2;
  uint8_t *w = static_cast<uint8_t *>(workspace);
  uint8_t *R[2] = {nullptr, nullptr};
  if (N > 1) { R[0] = w; w += up256(2 * kv_bytes); R[1] = w; w += up256(2 * kv_bytes); }
  MergeArgs m;
  memset(&m, 0, sizeof(m));
  m.n = N;
  const ReserveGuard reserve(ring_sm_reserve(N));
  if (N > 1) {
    FA_CUDA(cudaEventRecord(c->ev_in, st));
    FA_CUDA(cudaStreamWaitEvent(c->cs, c->ev_in, 0));
  }
  for (int t = 0; t < N; ++t) {
    const uint8_t *ck = t == 0 ? static_cast<const uint8_t *>(k) : R[(t - 1) & 1];
    const uint8_t *cv = t == 0 ? static_cast<const uint8_t *>(v) : R[(t - 1) & 1] + kv_bytes;
    if (t + 1 < N) {
      // the exchange for step t + 1 runs under step t's kernels; its receive buffer was read by step t - 1
      if (t >= 1) FA_CUDA(cudaStreamWaitEvent(c->cs, c->ev_done[t - 1], 0));
      FA_CALL(exchange(fn, c, ck, R[t & 1], kv_bytes, cv, R[t & 1] + kv_bytes, kv_bytes));
      FA_CUDA(cudaEventRecord(c->ev_xchg[t + 1], c->cs));
    }
    const int src = (c->rank - t + N) % N;
    const Step sp = ring_step(is_causal, c->rank, src, s_local);
    const int rows = sp.q1 - sp.q0, krows = sp.k1 - sp.k0;
    uint8_t *o_t = w; w += up256((size_t)b * s_local * h * d * 2);
    float *l_t = reinterpret_cast<float *>(w); w += up256((size_t)b * h * s_local * 4);
    m.o[t] = o_t; m.lse[t] = l_t; m.q0[t] = sp.q0; m.rows[t] = rows;
    for (int bi = 0; bi < b; ++bi) {
      const size_t qoff = (((size_t)bi * s_local + sp.q0) * h) * d * 2, koff = (((size_t)bi * s_local + sp.k0) * hk) * d * 2;
      if (!flash_attn_b200_fwd_shard(bi, static_cast<const uint8_t *>(q) + qoff, ck + koff, cv + koff, nullptr,
                                     o_t + (size_t)bi * rows * h * d * 2, nullptr, l_t + (size_t)bi * h * rows, 1, rows, krows,
                                     (rows + 127) / 128 * 128, (krows + 127) / 128 * 128, h, hk, d, (d + 31) / 32 * 32,
                                     p_dropout, softmax_scale, softmax_scale != 0.f ? 1.f / softmax_scale : 0.f, sp.causal,
                                     false, is_bf16, stream, seed, step_offset(offset, c->rank, src, N), nullptr, nullptr))
        return false;
    }
    if (t + 1 < N) {
      FA_CUDA(cudaEventRecord(c->ev_done[t], st));
      FA_CUDA(cudaStreamWaitEvent(st, c->ev_xchg[t + 1], 0));
    }
  }
  const long long warps = (long long)b * s_local * h;
  if (warps >= (1ll << 31)) return failf("%s: batch * s_local * heads must stay below 2^31", fn);
  const int blocks = (int)((warps * 32 + 255) / 256);
  if (is_bf16)
    ring_merge_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(m, static_cast<__nv_bfloat16 *>(out), static_cast<float *>(lse),
                                                             b, s_local, h, d);
  else
    ring_merge_kernel<__half><<<blocks, 256, 0, st>>>(m, static_cast<__half *>(out), static_cast<float *>(lse), b, s_local, h, d);
  FA_CUDA(cudaGetLastError());
  return true;
}

bool flash_attn_b200_ring_bwd_dropout(fa_ring_comm_t *c, const void *dout, const void *q, const void *k, const void *v,
                                      const void *out, const void *lse, void *dq, void *dk, void *dv, void *workspace,
                                      size_t workspace_bytes, int b, int s_local, int h, int hk, int d, float softmax_scale,
                                      bool is_causal, bool is_bf16, fa_stream_t stream, float p_dropout, uint64_t seed,
                                      uint64_t offset) {
  const char *fn = "flash_attn_b200_ring_bwd";
  const RingShape r{b, s_local, h, hk, d, is_bf16};
  FA_CALL(check_shape(fn, c, r, is_causal));
  if (!(p_dropout >= 0.f && p_dropout <= 1.f)) return failf("%s: p_dropout must be in [0, 1]", fn);
  if (!dout || !q || !k || !v || !out || !lse || !dq || !dk || !dv || !workspace) return failf("%s: NULL pointer", fn);
  const int N = c->nranks;
  if (workspace_bytes < flash_attn_b200_ring_bwd_workspace_bytes(N, b, s_local, h, hk, d))
    return failf("%s: workspace too small (flash_attn_b200_ring_bwd_workspace_bytes)", fn);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t s_r = ((size_t)s_local + 127) / 128 * 128, d_r = ((size_t)d + 31) / 32 * 32;
  const size_t kv_elems = (size_t)b * s_local * hk * d, kv_bytes = kv_elems * 2, dkv_bytes = kv_elems * 4;
  uint8_t *w = static_cast<uint8_t *>(workspace);
  uint8_t *R[2] = {nullptr, nullptr};
  if (N > 1) { R[0] = w; w += up256(2 * kv_bytes); R[1] = w; w += up256(2 * kv_bytes); }
  uint8_t *D[2] = {w, nullptr};
  w += up256(2 * dkv_bytes);
  if (N > 1) { D[1] = w; w += up256(2 * dkv_bytes); }
  float *dq_acc = reinterpret_cast<float *>(w); w += up256((size_t)b * s_local * h * d * 4);
  uint8_t *dq_t = w; w += up256((size_t)s_local * h * d * 2);
  uint8_t *dk_t = w; w += up256((size_t)s_local * hk * d * 2);
  uint8_t *dv_t = w; w += up256((size_t)s_local * hk * d * 2);
  float *lse_c = reinterpret_cast<float *>(w); w += up256((size_t)h * s_local * 4);
  float *softmax_d = reinterpret_cast<float *>(w); w += up256((size_t)h * s_r * 4);
  float *dq_accum = reinterpret_cast<float *>(w);

  const ReserveGuard reserve(ring_sm_reserve(N));
  FA_CUDA(cudaMemsetAsync(dq_acc, 0, (size_t)b * s_local * h * d * 4, st));
  FA_CUDA(cudaMemsetAsync(D[0], 0, 2 * dkv_bytes, st));
  if (N > 1) {
    FA_CUDA(cudaEventRecord(c->ev_in, st));
    FA_CUDA(cudaStreamWaitEvent(c->cs, c->ev_in, 0));
  }
  int cur = 0;                                                   // D[cur]: the sums of the chunk this rank holds now
  for (int t = 0; t < N; ++t) {
    const uint8_t *ck = t == 0 ? static_cast<const uint8_t *>(k) : R[(t - 1) & 1];
    const uint8_t *cv = t == 0 ? static_cast<const uint8_t *>(v) : R[(t - 1) & 1] + kv_bytes;
    if (t + 1 < N) {
      if (t >= 1) FA_CUDA(cudaStreamWaitEvent(c->cs, c->ev_done[t - 1], 0));
      FA_CALL(exchange(fn, c, ck, R[t & 1], kv_bytes, cv, R[t & 1] + kv_bytes, kv_bytes));
      FA_CUDA(cudaEventRecord(c->ev_xchg[t + 1], c->cs));
    }
    const int src = (c->rank - t + N) % N;
    const Step sp = ring_step(is_causal, c->rank, src, s_local);
    const int rows = sp.q1 - sp.q0, krows = sp.k1 - sp.k0;
    if (t > 0) cur ^= 1;       // D[cur]: the sums of the chunk held now, on their way in from the previous rank
    bool sums_here = t == 0;
    for (int bi = 0; bi < b; ++bi) {
      const size_t qoff = (((size_t)bi * s_local + sp.q0) * h) * d * 2, koff = (((size_t)bi * s_local + sp.k0) * hk) * d * 2;
      FA_CUDA(cudaMemcpy2DAsync(lse_c, (size_t)rows * 4, static_cast<const float *>(lse) + (size_t)bi * h * s_local + sp.q0,
                                (size_t)s_local * 4, (size_t)rows * 4, h, cudaMemcpyDeviceToDevice, st));
      if (!flash_attn_b200_bwd_shard(bi, static_cast<const uint8_t *>(dout) + qoff, static_cast<const uint8_t *>(q) + qoff,
                                     ck + koff, cv + koff, static_cast<const uint8_t *>(out) + qoff, softmax_d, lse_c, nullptr,
                                     dq_t, dk_t, dv_t, dq_accum, 1, rows, krows, (rows + 127) / 128 * 128,
                                     (krows + 127) / 128 * 128, h, hk, d, (int)d_r, p_dropout, softmax_scale,
                                     softmax_scale != 0.f ? 1.f / softmax_scale : 0.f, sp.causal, is_bf16, 0, stream, seed,
                                     step_offset(offset, c->rank, src, N), nullptr, nullptr))
        return false;
      FA_CALL(flash_attn_b200_cp_accumulate(dq_acc + ((size_t)bi * s_local + sp.q0) * h * d, dq_t, (int64_t)rows * h * d,
                                            is_bf16, stream));
      if (!sums_here) {
        // the partial sums arrive from the previous rank (sent there after its step t - 1).  Waited for HERE, not before
        // the step's kernels: only the two additions below need them, the 134 MB transfer runs under flash_attn_bwd
        FA_CUDA(cudaStreamWaitEvent(st, c->ev_dkv[t - 1], 0));
        sums_here = true;
      }
      float *dk_sum = reinterpret_cast<float *>(D[cur]) + ((size_t)bi * s_local + sp.k0) * hk * d;
      FA_CALL(flash_attn_b200_cp_accumulate(dk_sum, dk_t, (int64_t)krows * hk * d, is_bf16, stream));
      FA_CALL(flash_attn_b200_cp_accumulate(dk_sum + kv_elems, dv_t, (int64_t)krows * hk * d, is_bf16, stream));
    }
    if (N > 1) {
      FA_CUDA(cudaEventRecord(c->ev_done[t], st));               // kernels of step t done AND sums updated
      // pass the sums on; after the last step this hop is the way home
      FA_CUDA(cudaStreamWaitEvent(c->cs, c->ev_done[t], 0));
      FA_CALL(exchange(fn, c, D[cur], D[cur ^ 1], 2 * dkv_bytes, nullptr, nullptr, 0));
      FA_CUDA(cudaEventRecord(c->ev_dkv[t], c->cs));
      if (t + 1 < N) FA_CUDA(cudaStreamWaitEvent(st, c->ev_xchg[t + 1], 0));
    }
  }
  if (N > 1) {
    FA_CUDA(cudaStreamWaitEvent(st, c->ev_dkv[N - 1], 0));
    cur ^= 1;
  }
  FA_CALL(flash_attn_b200_cp_store(dq, dq_acc, (int64_t)b * s_local * h * d, is_bf16, stream));
  FA_CALL(flash_attn_b200_cp_store(dk, D[cur], (int64_t)kv_elems, is_bf16, stream));
  FA_CALL(flash_attn_b200_cp_store(dv, reinterpret_cast<float *>(D[cur]) + kv_elems, (int64_t)kv_elems, is_bf16, stream));
  return true;
}

bool flash_attn_b200_ring_fwd(fa_ring_comm_t *c, const void *q, const void *k, const void *v, void *out, void *lse,
                              void *workspace, size_t workspace_bytes, int b, int s_local, int h, int hk, int d,
                              float softmax_scale, bool is_causal, bool is_bf16, fa_stream_t stream) {
  return flash_attn_b200_ring_fwd_dropout(c, q, k, v, out, lse, workspace, workspace_bytes, b, s_local, h, hk, d, softmax_scale,
                                          is_causal, is_bf16, stream, 0.f, 0, 0);
}

bool flash_attn_b200_ring_bwd(fa_ring_comm_t *c, const void *dout, const void *q, const void *k, const void *v,
                              const void *out, const void *lse, void *dq, void *dk, void *dv, void *workspace,
                              size_t workspace_bytes, int b, int s_local, int h, int hk, int d, float softmax_scale,
                              bool is_causal, bool is_bf16, fa_stream_t stream) {
  return flash_attn_b200_ring_bwd_dropout(c, dout, q, k, v, out, lse, dq, dk, dv, workspace, workspace_bytes, b, s_local, h, hk,
                                          d, softmax_scale, is_causal, is_bf16, stream, 0.f, 0, 0);
}

}  // extern "C"
