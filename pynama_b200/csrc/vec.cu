// FP64 vector kernels (the PETSc Vec calls of the hot path).  Grid-stride, two-pass deterministic reductions.
#include <mutex>

#include "common.cuh"

namespace {
constexpr int BLK = 256;

__global__ void set_kernel(int64_t n, double a, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) x[i] = a;
}
__global__ void scale_kernel(int64_t n, double a, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) x[i] *= a;
}
__global__ void axpy_kernel(int64_t n, double a, const double* __restrict__ x, double* __restrict__ y) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) y[i] += a * x[i];
}
__global__ void aypx_kernel(int64_t n, double a, const double* __restrict__ x, double* __restrict__ y) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) y[i] = x[i] + a * y[i];
}
__global__ void waxpy_kernel(int64_t n, double a, const double* __restrict__ x, const double* __restrict__ y,
                             double* __restrict__ w) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) w[i] = a * x[i] + y[i];
}
struct MaxpyArgs {
  const double* x[8];
  double c[8];
  int m;
};
// y0, y and the x_j may alias (in-place updates X = X + sum c_j K_j, r = b - r): no __restrict__ here
__global__ void maxpy_kernel(int64_t n, const __grid_constant__ MaxpyArgs a, const double* y0, double* y) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    double s = y0 ? y0[i] : 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (j < a.m) s += a.c[j] * a.x[j][i];
    y[i] = s;
  }
}
// kind 0: dot(x,y), 1: sum x^2, 2: max |x|
template <int KIND>
__global__ void reduce1_kernel(int64_t n, const double* __restrict__ x, const double* __restrict__ y,
                               double* __restrict__ part) {
  __shared__ double sm[BLK / 32];
  double s = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    if (KIND == 0) s += x[i] * y[i];
    else if (KIND == 1) s += x[i] * x[i];
    else s = fmax(s, fabs(x[i]));
  }
  s = (KIND == 2) ? pyn::block_max<BLK>(s, sm) : pyn::block_sum<BLK>(s, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}
template <int KIND>
__global__ void reduce2_kernel(int nparts, const double* __restrict__ part, double* __restrict__ out) {
  __shared__ double sm[BLK / 32];
  double s = 0.0;
  for (int i = threadIdx.x; i < nparts; i += BLK) s = (KIND == 2) ? fmax(s, part[i]) : s + part[i];
  s = (KIND == 2) ? pyn::block_max<BLK>(s, sm) : pyn::block_sum<BLK>(s, sm);
  if (threadIdx.x == 0) out[0] = s;
}
__global__ void wrms1_kernel(int64_t n, const double* __restrict__ e, const double* __restrict__ y,
                             const double* __restrict__ yn, double atol, double rtol, double* __restrict__ part) {
  __shared__ double sm[BLK / 32];
  double s = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const double tol = atol + rtol * fmax(fabs(y[i]), fabs(yn[i]));
    const double q = e[i] / tol;
    s += q * q;
  }
  s = pyn::block_sum<BLK>(s, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}
__global__ void scatter_set_kernel(int64_t n, const int64_t* __restrict__ idx, const double* __restrict__ v,
                                   double cval, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK)
    x[idx[i]] = v ? v[i] : cval;
}
__global__ void vtensv_kernel(int64_t nnodes, const double* __restrict__ vel, double* __restrict__ out) {
  for (int64_t k = blockIdx.x * (int64_t)BLK + threadIdx.x; k < nnodes; k += (int64_t)gridDim.x * BLK) {
    const double vx = vel[2 * k], vy = vel[2 * k + 1];
    out[3 * k] = vx * vx;
    out[3 * k + 1] = vx * vy;
    out[3 * k + 2] = vy * vy;
  }
}
__global__ void taylor_green_kernel(int64_t n, const int64_t* __restrict__ idx, const double* __restrict__ xy, double nu,
                                    double t, double* __restrict__ vel, double* __restrict__ vort) {
  // custom_func.py:173-193 with Lx = Ly = Uref = 1
  const double pi = 3.141592653589793;
  const double expon = exp(-4.0 * (pi * pi) * nu * t * 2.0);
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const int64_t k = idx ? idx[i] : i;
    const double x_ = 2.0 * pi * xy[2 * k], y_ = 2.0 * pi * xy[2 * k + 1];
    double sx, cx, sy, cy;
    sincos(x_, &sx, &cx);
    sincos(y_, &sy, &cy);
    if (vel) {
      vel[2 * k] = cx * sy * expon;
      vel[2 * k + 1] = -sx * cy * expon;
    }
    if (vort) vort[k] = -2.0 * pi * 2.0 * cx * cy * expon;
  }
}

struct Scratch {
  double* part = nullptr;
  double* out = nullptr;
  double* h_out = nullptr;
  int cap = 0;
};
thread_local Scratch g_s;
int ensure_scratch() {
  if (g_s.part) return PYN_OK;
  g_s.cap = pyn::sm_count() * 8;
  PYN_CHECK_CUDA(cudaMalloc(&g_s.part, sizeof(double) * g_s.cap));
  PYN_CHECK_CUDA(cudaMalloc(&g_s.out, sizeof(double)));
  PYN_CHECK_CUDA(cudaMallocHost(&g_s.h_out, sizeof(double)));
  return PYN_OK;
}
template <int KIND>
int reduce(int64_t n, const double* x, const double* y, double* h_out, cudaStream_t st) {
  int rc = ensure_scratch();
  if (rc) return rc;
  int grid = pyn::grid_for(n, BLK, 8);
  reduce1_kernel<KIND><<<grid, BLK, 0, st>>>(n, x, y, g_s.part);
  PYN_CHECK_LAUNCH();
  reduce2_kernel<KIND><<<1, BLK, 0, st>>>(grid, g_s.part, g_s.out);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaMemcpyAsync(g_s.h_out, g_s.out, sizeof(double), cudaMemcpyDeviceToHost, st));
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  *h_out = *g_s.h_out;
  return PYN_OK;
}
}  // namespace

#define VEC_LAUNCH(kernel, n, ...)                                                              \
  do {                                                                                          \
    if ((n) <= 0) return PYN_OK;                                                                \
    kernel<<<pyn::grid_for((n), BLK, 8), BLK, 0, pyn::as_stream(stream)>>>(__VA_ARGS__);        \
    PYN_CHECK_LAUNCH();                                                                         \
    return PYN_OK;                                                                              \
  } while (0)

extern "C" {
int pyn_vec_set(int64_t n, double a, double* x, void* stream) { VEC_LAUNCH(set_kernel, n, n, a, x); }
int pyn_vec_scale(int64_t n, double a, double* x, void* stream) { VEC_LAUNCH(scale_kernel, n, n, a, x); }
int pyn_vec_axpy(int64_t n, double a, const double* x, double* y, void* stream) { VEC_LAUNCH(axpy_kernel, n, n, a, x, y); }
int pyn_vec_aypx(int64_t n, double a, const double* x, double* y, void* stream) { VEC_LAUNCH(aypx_kernel, n, n, a, x, y); }
int pyn_vec_waxpy(int64_t n, double a, const double* x, const double* y, double* w, void* stream) {
  VEC_LAUNCH(waxpy_kernel, n, n, a, x, y, w);
}
int pyn_vec_maxpy(int64_t n, int m, const double* h_coef, const double* const* h_xs, const double* y0, double* y,
                  void* stream) {
  PYN_REQUIRE(m >= 0 && m <= 8 && y, "maxpy supports at most 8 vectors");
  MaxpyArgs a;
  a.m = m;
  for (int j = 0; j < 8; ++j) {
    a.x[j] = j < m ? h_xs[j] : nullptr;
    a.c[j] = j < m ? h_coef[j] : 0.0;
  }
  VEC_LAUNCH(maxpy_kernel, n, n, a, y0, y);
}
int pyn_vec_dot(int64_t n, const double* x, const double* y, double* h_out, void* stream) {
  PYN_REQUIRE(n > 0 && x && y && h_out, "bad arguments");
  return reduce<0>(n, x, y, h_out, pyn::as_stream(stream));
}
int pyn_vec_norm(int64_t n, const double* x, int kind, double* h_out, void* stream) {
  PYN_REQUIRE(n > 0 && x && h_out && (kind == 2 || kind == 3), "bad arguments");
  if (kind == 3) return reduce<2>(n, x, nullptr, h_out, pyn::as_stream(stream));
  int rc = reduce<1>(n, x, nullptr, h_out, pyn::as_stream(stream));
  if (rc == PYN_OK) *h_out = sqrt(*h_out);
  return rc;
}
int pyn_vec_scatter_set(int64_t n, const int64_t* idx, const double* vals, double* x, void* stream) {
  PYN_REQUIRE(n >= 0 && (n == 0 || (idx && vals && x)), "bad arguments");
  VEC_LAUNCH(scatter_set_kernel, n, n, idx, vals, 0.0, x);
}
int pyn_vec_scatter_const(int64_t n, const int64_t* idx, double val, double* x, void* stream) {
  PYN_REQUIRE(n >= 0 && (n == 0 || (idx && x)), "bad arguments");
  VEC_LAUNCH(scatter_set_kernel, n, n, idx, nullptr, val, x);
}
int pyn_vtensv(int64_t nnodes, const double* vel, double* out, void* stream) {
  PYN_REQUIRE(vel && out, "bad arguments");
  VEC_LAUNCH(vtensv_kernel, nnodes, nnodes, vel, out);
}
int pyn_taylor_green(int64_t n, const int64_t* idx, const double* xy, double nu, double t, double* vel, double* vort,
                     void* stream) {
  PYN_REQUIRE(xy, "bad arguments");
  VEC_LAUNCH(taylor_green_kernel, n, n, idx, xy, nu, t, vel, vort);
}
int pyn_vec_wrms(int64_t n, const double* e, const double* y, const double* yn, double atol, double rtol, double* h_out,
                 void* stream) {
  PYN_REQUIRE(n > 0 && e && y && yn && h_out, "bad arguments");
  int rc = ensure_scratch();
  if (rc) return rc;
  cudaStream_t st = pyn::as_stream(stream);
  int grid = pyn::grid_for(n, BLK, 8);
  wrms1_kernel<<<grid, BLK, 0, st>>>(n, e, y, yn, atol, rtol, g_s.part);
  PYN_CHECK_LAUNCH();
  reduce2_kernel<1><<<1, BLK, 0, st>>>(grid, g_s.part, g_s.out);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaMemcpyAsync(g_s.h_out, g_s.out, sizeof(double), cudaMemcpyDeviceToHost, st));
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  *h_out = sqrt(*g_s.h_out / (double)n);
  return PYN_OK;
}
}  // extern "C"

// ---- pyn_vec_dot_noalloc: the dot product for callers that must not allocate ----------------------------------------
// Same two kernels, same grid and same order of additions as pyn_vec_dot (bit-identical result), but the per-CTA partial
// sums go to a file-scope device array instead of cudaMalloc'ed scratch and the result is left in the caller's DEVICE
// buffer: no cudaMalloc / cudaMallocHost / cudaFree, no synchronisation, nothing but two launches on the stream.
// Every GPU that loads the library holds the 16 KiB array until exit.  All calls share it: nothing in it has to start at
// zero (a call overwrites the grid entries it then reads), but two calls that overlap on the GPU (different streams)
// race on it.
static __device__ __align__(256) double g_noalloc_part[PYN_VEC_DOT_NOALLOC_MAX_PARTS];

namespace {
constexpr int NOALLOC_MAX_DEV = 64;
struct NoallocDev {
  double* part;   // device address of g_noalloc_part on this device (the symbol itself is a host-side stand-in)
  int sms;
};
NoallocDev g_noalloc_dev[NOALLOC_MAX_DEV];
std::mutex g_noalloc_mu;

// once per device, with that device current
int noalloc_setup(NoallocDev** out) {
  int dev = -1;
  PYN_CHECK_CUDA(cudaGetDevice(&dev));
  PYN_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEV, "pyn_vec_dot_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  NoallocDev& d = g_noalloc_dev[dev];
  if (!d.part) {
    double* ptr = nullptr;
    int sms = 0;
    PYN_CHECK_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noalloc_part));
    PYN_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    PYN_REQUIRE(ptr && sms > 0, "pyn_vec_dot_noalloc: no device address for the scratch array");
    d.sms = sms;
    d.part = ptr;
  }
  *out = &d;
  return PYN_OK;
}
}  // namespace

extern "C" int pyn_vec_dot_noalloc(int64_t n, const double* x, const double* y, double* d_out, void* stream) {
  PYN_REQUIRE(n > 0 && x && y && d_out, "bad arguments");
  if (n > PYN_VEC_DOT_NOALLOC_MAX_N) {
    pyn::set_error("pyn_vec_dot_noalloc: n = %lld exceeds the maximum of %lld entries", (long long)n,
                   (long long)PYN_VEC_DOT_NOALLOC_MAX_N);
    return PYN_ERR_ARG;
  }
  NoallocDev* d = nullptr;
  const int rc = noalloc_setup(&d);
  if (rc) return rc;
  int64_t g = (n + BLK - 1) / BLK;                       // the grid of pyn_vec_dot (grid_for(n, BLK, 8))
  if (g > (int64_t)d->sms * 8) g = (int64_t)d->sms * 8;
  if (g > PYN_VEC_DOT_NOALLOC_MAX_PARTS) {
    pyn::set_error("pyn_vec_dot_noalloc: %lld partial sums exceed the scratch array of %d", (long long)g,
                   PYN_VEC_DOT_NOALLOC_MAX_PARTS);
    return PYN_ERR_ARG;
  }
  cudaStream_t st = pyn::as_stream(stream);
  reduce1_kernel<0><<<(int)g, BLK, 0, st>>>(n, x, y, d->part);
  PYN_CHECK_LAUNCH();
  reduce2_kernel<0><<<1, BLK, 0, st>>>((int)g, d->part, d_out);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}
