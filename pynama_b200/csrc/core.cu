// Error reporting, launch accounting, device queries.
#include <stdarg.h>

#include <atomic>

#include "common.cuh"

namespace pyn {
static thread_local char g_err[1024] = "";
static std::atomic<int64_t> g_launches{0};
static int g_sm_count = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
  if (g_sm_count == 0) {
    int dev = 0, n = 0;
    // temporaries come from the stream-ordered allocator (cudaMallocAsync); keep what it frees in the pool instead of
    // handing it back to the driver at every synchronisation (unmapping and remapping costs milliseconds per step)
    cudaMemPool_t pool;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
      unsigned long long keep = ~0ull;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      g_sm_count = n;
    else
      g_sm_count = 148;
  }
  return g_sm_count;
}
}  // namespace pyn

extern "C" {
const char* pyn_last_error(void) { return pyn::g_err; }
int pyn_version(void) { return 100; }
int64_t pyn_launch_count(void) { return pyn::g_launches.load(); }

int pyn_device_info(int* sm, int* major, int* minor, int64_t* smem_optin) {
  int dev = 0;
  PYN_CHECK_CUDA(cudaGetDevice(&dev));
  int v = 0;
  PYN_CHECK_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev));
  if (sm) *sm = v;
  PYN_CHECK_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, dev));
  if (major) *major = v;
  PYN_CHECK_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, dev));
  if (minor) *minor = v;
  PYN_CHECK_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  if (smem_optin) *smem_optin = v;
  return PYN_OK;
}
}
