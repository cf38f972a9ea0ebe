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

// acc += blk, 8 elements per thread and trip (one 16-byte load of the 16-bit block, two 16-byte read-modify-writes of the
// fp32 accumulator); n8 = n / 8, the tail (n % 8 elements) is handled by the last threads one element at a time.
template <typename T>
__global__ void __launch_bounds__(256)
cp_accumulate_kernel(float *__restrict__ acc, const T *__restrict__ blk, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x, tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t n8 = n / 8;
  for (int64_t i = tid; i < n8; i += stride) {
    const uint4 v = reinterpret_cast<const uint4 *>(blk)[i];
    const T *e = reinterpret_cast<const T *>(&v);
    float4 a0 = reinterpret_cast<float4 *>(acc)[2 * i], a1 = reinterpret_cast<float4 *>(acc)[2 * i + 1];
    a0.x += to_f32<T>(e[0]); a0.y += to_f32<T>(e[1]); a0.z += to_f32<T>(e[2]); a0.w += to_f32<T>(e[3]);
    a1.x += to_f32<T>(e[4]); a1.y += to_f32<T>(e[5]); a1.z += to_f32<T>(e[6]); a1.w += to_f32<T>(e[7]);
    reinterpret_cast<float4 *>(acc)[2 * i] = a0;
    reinterpret_cast<float4 *>(acc)[2 * i + 1] = a1;
  }
  for (int64_t i = n8 * 8 + tid; i < n; i += stride) acc[i] += to_f32<T>(blk[i]);
}

template <typename T>
__global__ void __launch_bounds__(256)
cp_store_kernel(T *__restrict__ dst, const float *__restrict__ src, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x, tid = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int64_t n8 = n / 8;
  for (int64_t i = tid; i < n8; i += stride) {
    const float4 a0 = reinterpret_cast<const float4 *>(src)[2 * i], a1 = reinterpret_cast<const float4 *>(src)[2 * i + 1];
    T o[8] = {from_f32<T>(a0.x), from_f32<T>(a0.y), from_f32<T>(a0.z), from_f32<T>(a0.w),
              from_f32<T>(a1.x), from_f32<T>(a1.y), from_f32<T>(a1.z), from_f32<T>(a1.w)};
    reinterpret_cast<uint4 *>(dst)[i] = *reinterpret_cast<const uint4 *>(o);
  }
  for (int64_t i = n8 * 8 + tid; i < n; i += stride) dst[i] = from_f32<T>(src[i]);
}

// scalar versions for buffers that are not 16-byte aligned
template <typename T>
__global__ void __launch_bounds__(256)
cp_accumulate_scalar_kernel(float *__restrict__ acc, const T *__restrict__ blk, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += stride) acc[i] += to_f32<T>(blk[i]);
}

template <typename T>
__global__ void __launch_bounds__(256)
cp_store_scalar_kernel(T *__restrict__ dst, const float *__restrict__ src, int64_t n) {
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

namespace {
bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
}  // namespace

const char *launch_cp_accumulate(void *acc, const void *blk, int64_t n, bool is_bf16, cudaStream_t stream) {
  if (n <= 0) return nullptr;
  const bool vec = aligned16(acc) && aligned16(blk);
  const int grid = grid_for(vec ? (n + 7) / 8 : n);
  if (is_bf16) {
    if (vec) cp_accumulate_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((float *)acc, (const __nv_bfloat16 *)blk, n);
    else cp_accumulate_scalar_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((float *)acc, (const __nv_bfloat16 *)blk, n);
  } else {
    if (vec) cp_accumulate_kernel<__half><<<grid, 256, 0, stream>>>((float *)acc, (const __half *)blk, n);
    else cp_accumulate_scalar_kernel<__half><<<grid, 256, 0, stream>>>((float *)acc, (const __half *)blk, n);
  }
  return last_launch_error();
}

const char *launch_cp_store(void *dst, const void *src, int64_t n, bool is_bf16, cudaStream_t stream) {
  if (n <= 0) return nullptr;
  const bool vec = aligned16(dst) && aligned16(src);
  const int grid = grid_for(vec ? (n + 7) / 8 : n);
  if (is_bf16) {
    if (vec) cp_store_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((__nv_bfloat16 *)dst, (const float *)src, n);
    else cp_store_scalar_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((__nv_bfloat16 *)dst, (const float *)src, n);
  } else {
    if (vec) cp_store_kernel<__half><<<grid, 256, 0, stream>>>((__half *)dst, (const float *)src, n);
    else cp_store_scalar_kernel<__half><<<grid, 256, 0, stream>>>((__half *)dst, (const float *)src, n);
  }
  return last_launch_error();
}

}  // namespace fa
