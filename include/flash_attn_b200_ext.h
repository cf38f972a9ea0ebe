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

/* Number of kernel launches the last successful call on this thread enqueued. */
int flash_attn_b200_last_launch_count(void);

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
