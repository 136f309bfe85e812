// ORACLE SHIM (test infrastructure).  Stand-in for external/lce's <malloc_count/malloc_count.h>
// (missing submodule).  Same contract as the reference's own no-op stub, cmake/malloc_count_stub.c:15-27:
// every counter reads 0, so get_target_gap_idx_size (lz77_sss.hpp:117-122) is deterministic.
#pragma once
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif
static inline size_t malloc_count_current(void) { return 0; }
static inline size_t malloc_count_peak(void) { return 0; }
static inline void malloc_count_reset_peak(void) {}
#ifdef __cplusplus
}
#endif
