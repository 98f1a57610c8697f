// Shared host/device helpers for libhpe_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/hpe_b200.h"

namespace hpe {

// thread-local last-error string (hpe_last_error)
char* last_error_buf();
int   fail(int code, const char* fmt, ...);

#define HPE_CHECK_ARG(cond, ...)                                  \
  do { if (!(cond)) return ::hpe::fail(HPE_EINVAL, __VA_ARGS__); } while (0)

#define HPE_CUDA_OK(expr)                                                                     \
  do { cudaError_t _e = (expr);                                                               \
       if (_e != cudaSuccess)                                                                 \
         return ::hpe::fail(HPE_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                            __FILE__, __LINE__); } while (0)

inline int launch_status(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(HPE_ECUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
  return HPE_OK;
}

int num_sms();   // cached multiprocessor count of the current device

// ---- device helpers -----------------------------------------------------------------------------
// streaming 128-bit load: read-only path, do not allocate in L1 (each heatmap byte is used once)
__device__ __forceinline__ uint4 ld_stream_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_v4(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

}  // namespace hpe
