"""B200-native FlashAttention-2 behind the flash_attn.h C ABI.

The product is `lib/libflash_attn.so` (hand-written sm_100a CUDA, built by
`python -m flash_attention_hip_b200.build`).  `api` is a thin ctypes mirror of
the reference's operator interface for tests, benchmarks and Python callers.
"""
__version__ = "0.1"
