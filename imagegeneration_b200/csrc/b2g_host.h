// b2g_host.h — host-side helpers shared by the C-ABI translation units: error reporting and TMA descriptors.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200gan.h"

namespace b2g {

char* error_buffer();  // thread-local, 512 bytes (defined in b2g_api.cu)

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

#define B2G_CUDA_CHECK(expr)                                                                          \
  do {                                                                                                \
    cudaError_t e__ = (expr);                                                                         \
    if (e__ != cudaSuccess)                                                                           \
      return b2g::fail(B2G_E_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, \
                       __LINE__);                                                                     \
  } while (0)

#define B2G_LAUNCH_CHECK() B2G_CUDA_CHECK(cudaGetLastError())

// cuTensorMapEncodeTiled through the runtime's driver-entry-point lookup (no link-time libcuda dependency).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn();  // defined in b2g_api.cu; NULL when the driver is unavailable

// Tiled tensor map with 128-byte swizzle and zero fill for out-of-bounds elements.
// dims/box are innermost-first; strides_bytes has rank-1 entries (dimension 0 is contiguous).
inline int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                     const uint64_t* strides_bytes, const uint32_t* box,
                     CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(B2G_E_DRIVER, "cuTensorMapEncodeTiled unavailable (no CUDA driver)");
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], es[5];
  for (int i = 0; i < rank; ++i) {
    d[i] = dims[i];
    b[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
  CUresult r = fn(m, dt, (cuuint32_t)rank, const_cast<void*>(base), d, s, b, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    return fail(B2G_E_DRIVER,
                "cuTensorMapEncodeTiled failed (%d): rank %d base %p dims [%llu,%llu,%llu,%llu] box [%u,%u,%u,%u]",
                (int)r, rank, base, (unsigned long long)d[0], (unsigned long long)(rank > 1 ? d[1] : 0),
                (unsigned long long)(rank > 2 ? d[2] : 0), (unsigned long long)(rank > 3 ? d[3] : 0), b[0],
                rank > 1 ? b[1] : 0, rank > 2 ? b[2] : 0, rank > 3 ? b[3] : 0);
  }
  return B2G_OK;
}

int num_sms();     // SMs the launchers may fill: the device's count, or the b2g_set_sm_limit budget (b2g_api.cu)
int hw_num_sms();  // cached SM count of the current device

}  // namespace b2g
