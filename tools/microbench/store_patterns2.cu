// Micro-benchmark, part 2: what separates an SM store stream (5.9-6.3 TB/s in store_patterns.cu) from cudaMemset /
// torch fill_ (7.3-7.4 TB/s) on B200: grid shape, bytes per store instruction, or the stored values?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 --extended-lambda -o store_patterns2 store_patterns2.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void st32(double* p, double a, double b, double c, double d) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d) : "memory");
}
// one CTA per chunk of 256 x U x 16 B, U independent 16-byte stores per thread (the shape of torch's elementwise kernels)
template <int U>
__global__ void chunk16_kernel(double2* __restrict__ d, long long total2, double a, double b) {
  const double2 v = make_double2(a, b);
  const long long o = blockIdx.x * (long long)(256 * U) + threadIdx.x;
#pragma unroll
  for (int u = 0; u < U; ++u)
    if (o + u * 256 < total2) d[o + u * 256] = v;
}
// the same with 32-byte stores
template <int U>
__global__ void chunk32_kernel(double* __restrict__ d, long long total4, double a, double b) {
  const long long o = blockIdx.x * (long long)(256 * U) + threadIdx.x;
#pragma unroll
  for (int u = 0; u < U; ++u)
    if (o + u * 256 < total4) st32(d + 4 * (o + u * 256), a, b, a, b);
}
// persistent CTAs, contiguous batches, 32-byte stores
__global__ void persistent32_kernel(double* __restrict__ d, long long total4, int batch4, double a, double b) {
  const long long nb = (total4 + batch4 - 1) / batch4;
  for (long long bi = blockIdx.x; bi < nb; bi += gridDim.x) {
    const long long o = bi * (long long)batch4;
    const int n = (int)((total4 - o) < batch4 ? (total4 - o) : batch4);
    for (int i = threadIdx.x; i < n; i += blockDim.x) st32(d + 4 * (o + i), a, b, a, b);
  }
}
__global__ void persistent16_kernel(double2* __restrict__ d, long long total2, int batch2, double a, double b) {
  const double2 v = make_double2(a, b);
  const long long nb = (total2 + batch2 - 1) / batch2;
  for (long long bi = blockIdx.x; bi < nb; bi += gridDim.x) {
    const long long o = bi * (long long)batch2;
    const int n = (int)((total2 - o) < batch2 ? (total2 - o) : batch2);
#pragma unroll 4
    for (int i = threadIdx.x; i < n; i += blockDim.x) d[o + i] = v;
  }
}

template <class F>
static void run(const char* name, F launch, long long bytes) {
  cudaEvent_t s, e;
  cudaEventCreate(&s); cudaEventCreate(&e);
  for (int i = 0; i < 2; ++i) launch();
  cudaEventRecord(s);
  for (int i = 0; i < 5; ++i) launch();
  cudaEventRecord(e);
  cudaEventSynchronize(e);
  float ms; cudaEventElapsedTime(&ms, s, e); ms /= 5;
  cudaError_t err = cudaGetLastError();
  printf("{\"pattern\": \"%s\", \"ms\": %.4f, \"GBps\": %.1f%s}\n", name, ms, bytes / ms / 1e6, err ? ", \"error\": true" : "");
}

int main() {
  const long long total2 = 2412610692LL / 2, total4 = total2 / 2;
  const long long bytes = total2 * 16;
  double* d; if (cudaMalloc(&d, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  double2* d2 = (double2*)d;
  int sm = 148; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
  run("cudaMemset", [&] { cudaMemsetAsync(d, 0, bytes); }, bytes);
  run("chunk 16 B x4, constant value", [&] { chunk16_kernel<4><<<(unsigned)((total2 + 1023) / 1024), 256>>>(d2, total2, 1.5, 1.5); }, bytes);
  run("chunk 16 B x4, two values", [&] { chunk16_kernel<4><<<(unsigned)((total2 + 1023) / 1024), 256>>>(d2, total2, 1.5, 2.5); }, bytes);
  run("chunk 16 B x8, two values", [&] { chunk16_kernel<8><<<(unsigned)((total2 + 2047) / 2048), 256>>>(d2, total2, 1.5, 2.5); }, bytes);
  run("chunk 32 B x4, two values", [&] { chunk32_kernel<4><<<(unsigned)((total4 + 1023) / 1024), 256>>>(d, total4, 1.5, 2.5); }, bytes);
  run("chunk 32 B x2, two values", [&] { chunk32_kernel<2><<<(unsigned)((total4 + 511) / 512), 256>>>(d, total4, 1.5, 2.5); }, bytes);
  run("persistent 16 B, 204 kB batches, 4 CTAs/SM, constant value", [&] { persistent16_kernel<<<sm * 4, 256>>>(d2, total2, 12800, 1.5, 1.5); }, bytes);
  run("persistent 16 B, 204 kB batches, 4 CTAs/SM, two values", [&] { persistent16_kernel<<<sm * 4, 256>>>(d2, total2, 12800, 1.5, 2.5); }, bytes);
  run("persistent 32 B, 204 kB batches, 4 CTAs/SM, two values", [&] { persistent32_kernel<<<sm * 4, 256>>>(d, total4, 6400, 1.5, 2.5); }, bytes);
  run("persistent 32 B, 204 kB batches, 8 CTAs/SM, two values", [&] { persistent32_kernel<<<sm * 8, 256>>>(d, total4, 6400, 1.5, 2.5); }, bytes);
  run("persistent 32 B, 204 kB batches, 2 CTAs/SM, two values", [&] { persistent32_kernel<<<sm * 2, 256>>>(d, total4, 6400, 1.5, 2.5); }, bytes);
  cudaFree(d);
  return 0;
}
