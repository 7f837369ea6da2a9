// Shared helpers for the sm_100a kernels of the BOP pose-error hot path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/bop_b200.h"

// ---- error plumbing (api.cu owns the storage) -----------------------------------
int bop_set_error_cuda(cudaError_t e, const char* what, const char* file, int line);
int bop_set_error_msg(int code, const char* msg);
void bop_count_launch(int n = 1);
// Optional per-kernel CUDA-event timing (enabled by bop_profile_enable); slot ids below.
enum { BOP_PROF_VSD_FUSED = 0, BOP_PROF_VSD_PLAN = 1, BOP_PROF_MSSD = 2, BOP_PROF_MSPD = 3,
       BOP_PROF_K2 = 4, BOP_PROF_RENDER = 5, BOP_PROF_ADI = 6, BOP_PROF_MASK = 7, BOP_PROF_PNG = 8,
       BOP_PROF_SLOTS = 9 };
void bop_prof_begin(int slot, cudaStream_t s);
void bop_prof_end(int slot, cudaStream_t s);

#define BOP_CUDA(x)                                                        \
  do {                                                                     \
    cudaError_t _e = (x);                                                  \
    if (_e != cudaSuccess) return bop_set_error_cuda(_e, #x, __FILE__, __LINE__); \
  } while (0)

#define BOP_LAUNCH_CHECK()                                                 \
  do {                                                                     \
    cudaError_t _e = cudaGetLastError();                                   \
    if (_e != cudaSuccess) return bop_set_error_cuda(_e, "kernel launch", __FILE__, __LINE__); \
    bop_count_launch();                                                    \
  } while (0)

#define BOP_REQUIRE(cond, msg)                                             \
  do {                                                                     \
    if (!(cond)) return bop_set_error_msg(-1, msg);                        \
  } while (0)

// ---- device-side self checks ------------------------------------------------------
// Built with -DBOP_DEBUG_CHECKS (BOP_NVCC_FLAGS=-DBOP_DEBUG_CHECKS python -c "from bop_toolkit_b200
// import _lib; _lib.build()"), every index the kernels compute into shared-memory tiles, triangle
// lists, sort bins and images is range-checked and a violation traps (the launch then fails with
// an error instead of corrupting memory).  compute-sanitizer is not available on the GPU pool, so
// this is how the GPU test suite is run "with bounds checking" (profiles/r02_debug_checks.txt).
#ifdef BOP_DEBUG_CHECKS
#define BOP_DCHECK(cond)                                                                        \
  do {                                                                                          \
    if (!(cond)) {                                                                              \
      printf("BOP_DCHECK failed: %s (%s:%d) block %d thread %d\n", #cond, __FILE__, __LINE__,   \
             (int)blockIdx.x, (int)threadIdx.x);                                                \
      __trap();                                                                                 \
    }                                                                                           \
  } while (0)
#else
#define BOP_DCHECK(cond) do { } while (0)
#endif

// ---- device helpers -----------------------------------------------------------
#define BOP_FULL_MASK 0xffffffffu
#define BOP_INF_BITS_F32 0x7f800000u
#define BOP_INF_BITS_F64 0x7ff0000000000000ull

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(BOP_FULL_MASK, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(BOP_FULL_MASK, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum(int v) {
  return __reduce_add_sync(BOP_FULL_MASK, v);
}

// Streaming 128-bit load that does not allocate in L1 (depth images are read once).
__device__ __forceinline__ float4 ld_stream_f4(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint2 ld_stream_u2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
               : "=r"(r.x), "=r"(r.y)
               : "l"(p));
  return r;
}

// Streaming loads of the assembled-triangle list (written once, read once per row band).
__device__ __forceinline__ int4 ld_stream_i4(const int4* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ float ld_stream_f1(const float* p) {
  float r;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(r) : "l"(p));
  return r;
}

// 3x3 helpers on row-major doubles.
__device__ __forceinline__ void mat3_mul(const double* A, const double* B, double* C) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C[i * 3 + j] = A[i * 3 + 0] * B[0 * 3 + j] + A[i * 3 + 1] * B[1 * 3 + j] +
                     A[i * 3 + 2] * B[2 * 3 + j];
}
__device__ __forceinline__ void mat3_vec(const double* A, const double* v, double* o) {
#pragma unroll
  for (int i = 0; i < 3; ++i) o[i] = A[i * 3 + 0] * v[0] + A[i * 3 + 1] * v[1] + A[i * 3 + 2] * v[2];
}

