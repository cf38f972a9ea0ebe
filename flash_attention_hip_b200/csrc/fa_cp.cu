// fa_cp.cu -- small HBM-bound kernels for the ring context-parallel driver
// (BASELINE.json config 5; flash_attention_hip_b200/ring.py).  The attention
// math of every ring step is the ordinary flash_attn_fwd / flash_attn_bwd call
// (reference flash_attn.h:10-36, :68-99) on a (query slice, visiting K/V chunk)
// pair; these kernels only combine the per-step results:
//   cp_merge       (o_acc, lse_acc) <- log-sum-exp merge with one step's (out, lse)
//   cp_accumulate  fp32 accumulator += 16-bit block (dq / dk / dv partials)
//   cp_store       fp32 accumulator -> 16-bit tensor
// An "empty" row (no visible key so far) carries lse = +inf, the library's
// convention for fully masked rows.
#include "fa_common.cuh"

namespace fa {

namespace {

template <typename T>
__global__ void __launch_bounds__(256)
cp_merge_kernel(float *__restrict__ o_acc, float *__restrict__ lse_acc, const T *__restrict__ o_blk,
                const float *__restrict__ lse_blk, int rows, int h, int d, int64_t lse_acc_stride,
                int64_t lse_blk_stride) {
  const int warp = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (warp >= rows * h) return;
  const int row = warp / h, head = warp % h;
  float *la_p = lse_acc + head * lse_acc_stride + row;
  const float la = *la_p, lb = lse_blk[head * lse_blk_stride + row];
  if (lb == INFINITY) return;                          // this step saw no key for the row
  float wa, wb, ln;
  if (la == INFINITY) {                                // first contribution
    wa = 0.f; wb = 1.f; ln = lb;
  } else {
    const float m = fmaxf(la, lb);
    const float ea = __expf(la - m), eb = __expf(lb - m);
    ln = m + __logf(ea + eb);
    wa = ea / (ea + eb); wb = eb / (ea + eb);
  }
  float *oa = o_acc + ((size_t)row * h + head) * d;
  const T *ob = o_blk + ((size_t)row * h + head) * d;
  for (int c = lane; c < d; c += 32) oa[c] = wa * oa[c] + wb * to_f32<T>(ob[c]);
  __syncwarp();
  if (lane == 0) *la_p = ln;
}

template <typename T>
__global__ void __launch_bounds__(256)
cp_accumulate_kernel(float *__restrict__ acc, const T *__restrict__ blk, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride) acc[i] += to_f32<T>(blk[i]);
}

template <typename T>
__global__ void __launch_bounds__(256)
cp_store_kernel(T *__restrict__ dst, const float *__restrict__ src, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = from_f32<T>(src[i]);
}

int grid_for(int64_t n) {
  const int64_t blocks = (n + 255) / 256;
  return (int)(blocks < 148 * 16 ? (blocks > 0 ? blocks : 1) : 148 * 16);   // grid-stride, 16 CTAs per SM
}

const char *last_launch_error() {
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? nullptr : cudaGetErrorString(e);
}

}  // namespace

const char *launch_cp_merge(void *o_acc, void *lse_acc, const void *o_blk, const void *lse_blk, int rows, int h,
                            int d, int64_t lse_acc_stride, int64_t lse_blk_stride, bool is_bf16,
                            cudaStream_t stream) {
  if (rows <= 0) return nullptr;
  const int64_t threads = (int64_t)rows * h * 32;
  const int blocks = (int)((threads + 255) / 256);
  if (is_bf16)
    cp_merge_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(
        (float *)o_acc, (float *)lse_acc, (const __nv_bfloat16 *)o_blk, (const float *)lse_blk, rows, h, d,
        lse_acc_stride, lse_blk_stride);
  else
    cp_merge_kernel<__half><<<blocks, 256, 0, stream>>>((float *)o_acc, (float *)lse_acc, (const __half *)o_blk,
                                                        (const float *)lse_blk, rows, h, d, lse_acc_stride,
                                                        lse_blk_stride);
  return last_launch_error();
}

const char *launch_cp_accumulate(void *acc, const void *blk, int64_t n, bool is_bf16, cudaStream_t stream) {
  if (n <= 0) return nullptr;
  if (is_bf16)
    cp_accumulate_kernel<__nv_bfloat16><<<grid_for(n), 256, 0, stream>>>((float *)acc, (const __nv_bfloat16 *)blk, n);
  else
    cp_accumulate_kernel<__half><<<grid_for(n), 256, 0, stream>>>((float *)acc, (const __half *)blk, n);
  return last_launch_error();
}

const char *launch_cp_store(void *dst, const void *src, int64_t n, bool is_bf16, cudaStream_t stream) {
  if (n <= 0) return nullptr;
  if (is_bf16)
    cp_store_kernel<__nv_bfloat16><<<grid_for(n), 256, 0, stream>>>((__nv_bfloat16 *)dst, (const float *)src, n);
  else
    cp_store_kernel<__half><<<grid_for(n), 256, 0, stream>>>((__half *)dst, (const float *)src, n);
  return last_launch_error();
}

}  // namespace fa
