// Shared helpers of libpsob.so (error convention of include/psob.h).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>

#include "../../include/psob.h"

namespace psob {

extern thread_local char g_err[512];

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define PSOB_CUDA(call)                                                                         \
  do {                                                                                          \
    cudaError_t e__ = (call);                                                                   \
    if (e__ != cudaSuccess)                                                                     \
      return psob::fail((int)e__, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),      \
                        __FILE__, __LINE__);                                                    \
  } while (0)

#define PSOB_LAUNCH_CHECK(name)                                                                 \
  do {                                                                                          \
    cudaError_t e__ = cudaGetLastError();                                                       \
    if (e__ != cudaSuccess)                                                                     \
      return psob::fail((int)e__, "launch of %s failed: %s", name, cudaGetErrorString(e__));    \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// order-preserving float -> uint32 (NaN with sign 0 sorts above +inf); -0.0 is canonicalised to +0.0 so that the two
// zeros tie like they do under the reference's `<` (lowest index wins, src/pso/util_pso.py:139-143)
__host__ __device__ inline uint32_t order_key(float f) {
#ifdef __CUDA_ARCH__
  uint32_t b = __float_as_uint(f);
#else
  uint32_t b;
  memcpy(&b, &f, 4);
#endif
  if (b == 0x80000000u) b = 0u;
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

}  // namespace psob
