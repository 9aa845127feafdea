// Shared helpers for libpynama_b200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/pynama_b200.h"

namespace pyn {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define PYN_CHECK_CUDA(expr)                                                                   \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      pyn::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));    \
      return PYN_ERR_CUDA;                                                                     \
    }                                                                                          \
  } while (0)

#define PYN_CHECK_LAUNCH()                                                                     \
  do {                                                                                         \
    pyn::count_launch();                                                                       \
    cudaError_t _e = cudaGetLastError();                                                       \
    if (_e != cudaSuccess) {                                                                   \
      pyn::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return PYN_ERR_CUDA;                                                                     \
    }                                                                                          \
  } while (0)

#define PYN_REQUIRE(cond, msg)                                                                 \
  do {                                                                                         \
    if (!(cond)) {                                                                             \
      pyn::set_error("%s:%d: %s", __FILE__, __LINE__, msg);                                    \
      return PYN_ERR_ARG;                                                                      \
    }                                                                                          \
  } while (0)

int sm_count();

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

static inline int grid_for(int64_t work, int block, int max_waves = 8) {
  int64_t g = (work + block - 1) / block;
  int64_t cap = (int64_t)sm_count() * max_waves;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// ---- device helpers -------------------------------------------------------------------------------
#ifdef __CUDACC__
// streaming (read-once) loads: bypass L1 allocation so the gathered vector keeps the cache
__device__ __forceinline__ double ld_stream(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int ld_stream(const int* p) {
  int v;
  asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ double2 ld_stream2(const double* p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ int4 ld_stream4(const int* p) {
  int4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream(double* p, double v) {
  asm volatile("st.global.L1::no_allocate.f64 [%0], %1;" ::"l"(p), "d"(v));
}
__device__ __forceinline__ void st_stream(int* p, int v) {
  asm volatile("st.global.L1::no_allocate.s32 [%0], %1;" ::"l"(p), "r"(v));
}

template <int W>
__device__ __forceinline__ double subwarp_sum(double v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o, W);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) { return subwarp_sum<32>(v); }
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

// Deterministic block reduction (fixed tree): result valid in thread 0.
template <int BLOCK>
__device__ __forceinline__ double block_sum(double v, double* smem /* >= BLOCK/32 doubles */) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) smem[w] = v;
  __syncthreads();
  double r = 0.0;
  if (w == 0) {
    r = (lane < BLOCK / 32) ? smem[lane] : 0.0;
    r = warp_sum(r);
  }
  __syncthreads();
  return r;
}
template <int BLOCK>
__device__ __forceinline__ double block_max(double v, double* smem) {
  v = warp_max(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) smem[w] = v;
  __syncthreads();
  double r = 0.0;
  if (w == 0) {
    r = (lane < BLOCK / 32) ? smem[lane] : 0.0;
    r = warp_max(r);
  }
  __syncthreads();
  return r;
}
// Sum `n` partials in a fixed order with one warp-strided pass per block (every block gets the same bits).
__device__ __forceinline__ double reduce_partials(const double* __restrict__ part, int n) {
  // all threads of the calling warp participate
  double s = 0.0;
  for (int i = threadIdx.x & 31; i < n; i += 32) s += part[i];
  s = warp_sum(s);
  return __shfl_sync(0xffffffffu, s, 0);
}
#endif

}  // namespace pyn
