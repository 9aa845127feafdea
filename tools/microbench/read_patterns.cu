// Micro-benchmark, part 3: HBM READ bandwidth of a streaming reduction (the shape of the CSR-stream SpMV: 12 B per entry
// read once) under three work distributions: persistent CTAs striding over chunks in lock-step, persistent CTAs owning
// one contiguous block each, and one CTA per group of chunks dispatched in order.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 --extended-lambda -o read_patterns read_patterns.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double2 ldnc(const double2* p) {
  double2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ double chunk_sum(const double2* __restrict__ d, long long o, int n) {
  double s = 0.0;
#pragma unroll 8
  for (int i = threadIdx.x; i < n; i += blockDim.x) { const double2 v = ldnc(d + o + i); s += v.x + v.y; }
  return s;
}
// mode 0: strided persistent; 1: blocked persistent; 2: one CTA per `group` consecutive chunks (grid = nchunk / group)
__global__ void read_kernel(const double2* __restrict__ d, long long total2, int chunk2, int mode, int group, double* out) {
  const long long nch = (total2 + chunk2 - 1) / chunk2;
  double s = 0.0;
  long long b0, b1, st;
  if (mode == 0) { b0 = blockIdx.x; b1 = nch; st = gridDim.x; }
  else if (mode == 1) { const long long per = (nch + gridDim.x - 1) / gridDim.x; b0 = blockIdx.x * per; b1 = b0 + per < nch ? b0 + per : nch; st = 1; }
  else { b0 = blockIdx.x * (long long)group; b1 = b0 + group < nch ? b0 + group : nch; st = 1; }
  for (long long b = b0; b < b1; b += st) {
    const long long o = b * (long long)chunk2;
    s += chunk_sum(d, o, (int)((total2 - o) < chunk2 ? (total2 - o) : chunk2));
  }
  if (s == 123.456) out[blockIdx.x] = s;
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
  const long long total2 = 29000000000LL / 16;   // the K stream of cfg 4: 29 GB
  const long long bytes = total2 * 16;
  double2* d; if (cudaMalloc(&d, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(d, 0, bytes);
  double* out; cudaMalloc(&out, 8 << 20);
  int sm = 148; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
  const int chunk2 = 1536;   // 24 kB = 2048 entries x 12 B
  const long long nch = (total2 + chunk2 - 1) / chunk2;
  char name[160];
  for (int c : {4, 5, 8}) {
    snprintf(name, sizeof name, "persistent, strided chunks of 24 kB, %d CTAs/SM x 256", c);
    run(name, [&] { read_kernel<<<sm * c, 256>>>(d, total2, chunk2, 0, 0, out); }, bytes);
    snprintf(name, sizeof name, "persistent, one contiguous block per CTA, %d CTAs/SM x 256", c);
    run(name, [&] { read_kernel<<<sm * c, 256>>>(d, total2, chunk2, 1, 0, out); }, bytes);
  }
  for (int g : {1, 4, 16, 64}) {
    snprintf(name, sizeof name, "one CTA per %d consecutive chunks (grid %lld) x 256", g, (nch + g - 1) / g);
    run(name, [&] { read_kernel<<<(unsigned)((nch + g - 1) / g), 256>>>(d, total2, chunk2, 2, g, out); }, bytes);
  }
  cudaFree(d); cudaFree(out);
  return 0;
}
