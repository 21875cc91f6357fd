// Shared helpers for the sm_100a paged-KV kernels: error plumbing for the C ABI, dtype traits, vector
// load/store wrappers.  Nothing here allocates or synchronises.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/b200_kv.h"

namespace b200 {

// ---- error plumbing ---------------------------------------------------------------------------------
char* last_error_buf();  // thread-local, defined in b200_abi.cu
constexpr int kErrBufLen = 512;

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(last_error_buf(), kErrBufLen, fmt, ap);
  va_end(ap);
  return code;
}

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(static_cast<int>(e), "%s: %s", what, cudaGetErrorString(e));
  return 0;
}

inline int dtype_size(int dtype) {
  switch (dtype) {
    case B200_DTYPE_F16:
    case B200_DTYPE_BF16:
      return 2;
    case B200_DTYPE_F32:
      return 4;
    default:
      return 0;
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- device helpers ---------------------------------------------------------------------------------
template <typename T>
struct Cvt;
template <>
struct Cvt<float> {
  static __device__ __forceinline__ float to_f32(float x) { return x; }
  static __device__ __forceinline__ float from_f32(float x) { return x; }
};
template <>
struct Cvt<__half> {
  static __device__ __forceinline__ float to_f32(__half x) { return __half2float(x); }
  static __device__ __forceinline__ __half from_f32(float x) { return __float2half_rn(x); }
};
template <>
struct Cvt<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
  static __device__ __forceinline__ __nv_bfloat16 from_f32(float x) { return __float2bfloat16_rn(x); }
};

// streaming 16-byte global load / store: data that is touched once should not displace L1 lines
__device__ __forceinline__ int4 ld_stream16(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream16(void* p, const int4& v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kNumSMs = 148;  // B200

}  // namespace b200
