// Micro-benchmark: HBM write bandwidth of store streams shaped like asm_stream_kernel (persistent CTAs, each streaming
// a contiguous batch) against a plain fill, to see which access pattern the memory system rewards.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o store_patterns store_patterns.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

// every CTA (blockDim threads) streams batches of `batch2` double2, batch index = blockIdx + it * gridDim
__global__ void persistent_kernel(double2* __restrict__ d, long long total2, int batch2) {
  const double2 v = make_double2(1.5, 2.5);
  const long long nb = (total2 + batch2 - 1) / batch2;
  for (long long b = blockIdx.x; b < nb; b += gridDim.x) {
    const long long o = b * (long long)batch2;
    const int n = (int)((total2 - o) < batch2 ? (total2 - o) : batch2);
    for (int i = threadIdx.x; i < n; i += blockDim.x) d[o + i] = v;
  }
}
// the same, but every WARP owns its batch
__global__ void persistent_warp_kernel(double2* __restrict__ d, long long total2, int batch2) {
  const double2 v = make_double2(1.5, 2.5);
  const long long nb = (total2 + batch2 - 1) / batch2;
  const int lane = threadIdx.x & 31;
  const long long w = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5), nw = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long b = w; b < nb; b += nw) {
    const long long o = b * (long long)batch2;
    const int n = (int)((total2 - o) < batch2 ? (total2 - o) : batch2);
    for (int i = lane; i < n; i += 32) d[o + i] = v;
  }
}
// grid-stride fill: thread i writes i, i + T, ...
__global__ void gridstride_kernel(double2* __restrict__ d, long long total2) {
  const double2 v = make_double2(1.5, 2.5);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total2; i += (long long)gridDim.x * blockDim.x) d[i] = v;
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
  const long long total2 = 2412610692LL / 2;   // K of cfg 4 in double2
  const long long bytes = total2 * 16;
  double2* d; if (cudaMalloc(&d, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  int sm = 148; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, 0);
  run("cudaMemset", [&] { cudaMemsetAsync(d, 0, bytes); }, bytes);
  run("grid-stride fill, 148x8 CTAs x 256", [&] { gridstride_kernel<<<sm * 8, 256>>>(d, total2); }, bytes);
  const int batches[] = {512, 1600, 6400, 12800, 25600};   // double2 per batch: 8 kB .. 410 kB (256 interior nodes = 12800)
  const int ctas[] = {2, 4, 8};
  char name[128];
  for (int b : batches)
    for (int c : ctas) {
      snprintf(name, sizeof name, "persistent CTA batches of %d kB, %d CTAs/SM x 256", b * 16 / 1000, c);
      run(name, [&] { persistent_kernel<<<sm * c, 256>>>(d, total2, b); }, bytes);
    }
  for (int b : {400, 1600, 6400}) {
    snprintf(name, sizeof name, "persistent WARP batches of %d kB, 8 CTAs/SM x 256", b * 16 / 1000);
    run(name, [&] { persistent_warp_kernel<<<sm * 8, 256>>>(d, total2, b); }, bytes);
  }
  run("persistent CTA batches of 204 kB, 8 CTAs/SM x 512", [&] { persistent_kernel<<<sm * 4, 512>>>(d, total2, 12800); }, bytes);
  run("persistent CTA batches of 204 kB, 16 CTAs/SM x 128", [&] { persistent_kernel<<<sm * 16, 128>>>(d, total2, 12800); }, bytes);
  cudaFree(d);
  return 0;
}
