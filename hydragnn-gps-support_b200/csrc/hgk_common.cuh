// Shared host/device helpers for libhgk (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/hgk.h"

namespace hgk {

// Per-thread error text behind hgk_last_error().
char* error_buffer();
int fail(int status, const char* fmt, ...);

inline int check_launch(const char* what) {
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess)
    return fail(HGK_ERR_CUDA, "%s: %s", what, cudaGetErrorString(err));
  return HGK_OK;
}

#define HGK_REQUIRE(cond, ...)                                  \
  do {                                                          \
    if (!(cond)) return hgk::fail(HGK_ERR_INVALID_ARG, __VA_ARGS__); \
  } while (0)

constexpr int kNumSM = 148;  // B200

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
inline int64_t align_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }
inline int64_t imin64(int64_t a, int64_t b) { return a < b ? a : b; }

// Streaming 128-bit load that does not pollute L1 (edge tensors are read once).
__device__ __forceinline__ float4 ld_stream4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream1(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}
// Streaming 128-bit store (edge outputs are written once, consumed by a later kernel).
__device__ __forceinline__ void st_stream4(float* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p),
               "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// 256-bit accesses (sm_100: LDG.256 / STG.256): one full 32-byte sector per lane and instruction,
// which matters when the lanes of a warp walk different rows.  `p` must be 32-byte aligned.
__device__ __forceinline__ void st_global_v8(float* p, float4 a, float4 b) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z),
               "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w)
               : "memory");
}
__device__ __forceinline__ void ld_global_nc_v8(const float* p, float4& a, float4& b) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(p));
}

// IEEE multiply / add that the compiler may not contract into an FMA: the
// segmented sums must round exactly like "materialise product, then add".
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }

}  // namespace hgk
