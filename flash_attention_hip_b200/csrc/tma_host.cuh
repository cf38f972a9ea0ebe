// tma_host.cuh -- host-side TMA tensor-map construction.  cuTensorMapEncodeTiled
// is fetched through cudaGetDriverEntryPoint so the library links against the
// runtime only (no -lcuda).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace fa {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                  const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

// Tensor map over a [tokens, heads, d] 16-bit tensor viewed as 4-D
// (d, heads, rows, batches): dense layouts pass rows = seqlen, batches = b
// (out-of-range rows of one sequence then read as zero, never as the next
// batch entry's data); packed layouts pass rows = total tokens, batches = 1.
// Box = 64 elements of d (128 bytes, the swizzle span) x 1 head x box_rows rows.
inline const char *make_tmap_rows(CUtensorMap *m, const void *base, bool is_bf16, int d, int heads,
                                  uint64_t rows, uint64_t batches, int box_rows) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return "cuTensorMapEncodeTiled is unavailable (driver too old?)";
  const cuuint64_t dims[4] = {(cuuint64_t)d, (cuuint64_t)heads, rows ? rows : 1, batches ? batches : 1};
  const cuuint64_t strides[3] = {(cuuint64_t)d * 2, (cuuint64_t)heads * d * 2,
                                 (cuuint64_t)heads * d * 2 * (rows ? rows : 1)};
  const cuuint32_t box[4] = {64, 1, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                         const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? nullptr : "cuTensorMapEncodeTiled failed";
}

// The same 4-D view (d, heads, rows, batches) with explicit element strides for heads, rows and batches
// (each a multiple of 8 elements): lets a caller-side layout be read under another index order, e.g. the
// query heads of one GQA group as the ROWS of a tile (fa_tc_fwd.cu, decode-like calls).
inline const char *make_tmap_strided(CUtensorMap *m, const void *base, bool is_bf16, int d, uint64_t heads,
                                     uint64_t rows, uint64_t batches, uint64_t head_stride, uint64_t row_stride,
                                     uint64_t batch_stride, int box_rows) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return "cuTensorMapEncodeTiled is unavailable (driver too old?)";
  const cuuint64_t dims[4] = {(cuuint64_t)d, heads ? heads : 1, rows ? rows : 1, batches ? batches : 1};
  const cuuint64_t strides[3] = {head_stride * 2, row_stride * 2, batch_stride * 2};
  const cuuint32_t box[4] = {64, 1, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                         const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? nullptr : "cuTensorMapEncodeTiled failed";
}

// Additive bias [batches, heads, rows, cols] (16-bit, cols contiguous) as a 4-D map (cols, rows, heads, batches) with a
// box of box_cols x box_rows elements, NOT swizzled: a tile lands in shared memory as box_rows rows of box_cols * 2 bytes.
// cols * 2 bytes and the base must be multiples of 16 (the caller checks).
inline const char *make_tmap_bias(CUtensorMap *m, const void *base, bool is_bf16, uint64_t cols, uint64_t rows,
                                  uint64_t heads, uint64_t batches, int box_cols, int box_rows) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return "cuTensorMapEncodeTiled is unavailable (driver too old?)";
  const cuuint64_t dims[4] = {cols, rows, heads ? heads : 1, batches ? batches : 1};
  const cuuint64_t strides[3] = {cols * 2, rows * cols * 2, (heads ? heads : 1) * rows * cols * 2};
  const cuuint32_t bx[4] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  const CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                         const_cast<void *>(base), dims, strides, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? nullptr : "cuTensorMapEncodeTiled (bias) failed";
}

}  // namespace fa
