/* Two-line stand-in for <hip/hip_runtime.h> so that the reference header
 * (/root/reference/flash_attn.h:3) compiles verbatim on a CUDA-only box.
 * hipStream_t and cudaStream_t are both pointers to an opaque struct. */
#pragma once
typedef struct CUstream_st *hipStream_t;
