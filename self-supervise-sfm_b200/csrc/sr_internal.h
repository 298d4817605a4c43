// Host-side internals shared by the .cu translation units of libsailrecon_b200.so
// (error reporting, TMA descriptor encoding).  Not part of the public C-ABI.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/sailrecon_b200.h"

namespace sr {

// Records a message retrievable through sr_last_error(); returns `code`.
int fail(int code, const char* fmt, ...);

#define SR_CHECK_CUDA(expr)                                                                      \
  do {                                                                                           \
    cudaError_t _e = (expr);                                                                     \
    if (_e != cudaSuccess) return ::sr::fail(SR_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

#define SR_REQUIRE(cond, ...)                                  \
  do {                                                         \
    if (!(cond)) return ::sr::fail(SR_ERR_SHAPE, __VA_ARGS__); \
  } while (0)

// bf16 tensor map, SWIZZLE_128B, zero OOB fill.  dims/strides innermost first; strides in BYTES for dims 1..rank-1.
int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box);

// Profiling / experiment knobs (SR_ATTN_TRACE, SR_ATTN_POLY, SR_ATTN_DBG, SR_ATTN_LAG_NS, SR_ATTN_KERNEL, SR_ATTN_BWD_DBG) are read
// from the environment ONLY in -DSR_DEBUG builds; the shipped library ignores the environment entirely (some of the knobs
// skip arithmetic for timing experiments and would silently produce wrong results).
#ifdef SR_DEBUG
inline const char* dbg_env(const char* name) { return getenv(name); }
constexpr bool kDebugBuild = true;
#else
inline const char* dbg_env(const char*) { return nullptr; }
constexpr bool kDebugBuild = false;      // device-side experiment branches fold away
#endif

// Checks (once) that the current device is compute capability 10.x.
int require_sm100();

}  // namespace sr
