// y = A x with fused epilogues (see include/pynama_b200.h: pyn_spmv).
#include <stdlib.h>

#include "spmv.cuh"

namespace {

template <int LPR>
__global__ void __launch_bounds__(256)
spmv_kernel(int64_t nrows, const int64_t* __restrict__ ptr, const int* __restrict__ ind, const double* __restrict__ val,
            const double* __restrict__ x, double* __restrict__ y, int mode, double p0, double p1,
            const double* __restrict__ aux) {
  const int sub = threadIdx.x % LPR;
  const int64_t rows_per_pass = ((int64_t)gridDim.x * blockDim.x) / LPR;
  const int64_t first = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) / LPR;
  const int64_t passes = (nrows + rows_per_pass - 1) / rows_per_pass;
  for (int64_t it = 0; it < passes; ++it) {
    const int64_t row = first + it * rows_per_pass;
    const bool active = row < nrows;
    const double s = pyn::spmv_row<LPR>(ptr, ind, val, x, row, active, sub);
    if (active && sub == 0) {
      double out;
      switch (mode) {
        case PYN_SPMV_ADD: out = y[row] + s; break;
        case PYN_SPMV_SCALE: out = p0 * s; break;
        case PYN_SPMV_SRT_VTENSV: {
          const int64_t k = row / 3;
          const int c = (int)(row - 3 * k);
          const double vx = aux[2 * k], vy = aux[2 * k + 1];
          const double t = (c == 0) ? vx * vx : ((c == 1) ? vx * vy : vy * vy);
          out = p0 * s + (-p1) * t;
        } break;
        default: out = s;
      }
      y[row] = out;
    }
  }
}

template <int LPR>
int launch(int64_t nrows, const int64_t* ptr, const int* ind, const double* val, const double* x, double* y, int mode,
           double p0, double p1, const double* aux, cudaStream_t st) {
  const int block = 256;
  const int64_t need = (nrows * LPR + block - 1) / block;
  const int64_t cap = (int64_t)pyn::sm_count() * 16;
  const int grid = (int)(need < cap ? need : cap);
  spmv_kernel<LPR><<<grid, block, 0, st>>>(nrows, ptr, ind, val, x, y, mode, p0, p1, aux);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

// ---- CSR-stream kernel (planned) ------------------------------------------------------------------------------------
struct ModeEpi {
  double* y;
  int mode;
  double p0, p1;
  const double* aux;
  __device__ __forceinline__ void operator()(int64_t row, double s) const {
    double out;
    switch (mode) {
      case PYN_SPMV_ADD: out = y[row] + s; break;
      case PYN_SPMV_SCALE: out = p0 * s; break;
      case PYN_SPMV_SRT_VTENSV: {
        const int64_t k = row / 3;
        const int c = (int)(row - 3 * k);
        const double vx = aux[2 * k], vy = aux[2 * k + 1];
        const double t = (c == 0) ? vx * vx : ((c == 1) ? vx * vy : vy * vy);
        out = p0 * s + (-p1) * t;
      } break;
      default: out = s;
    }
    y[row] = out;
  }
};

__global__ void __launch_bounds__(pyn::SP_BLK, 6)
spmv_stream_kernel(pyn::SpmvPlanView P, const int* __restrict__ ind, const double* __restrict__ val,
                   const double* __restrict__ x, ModeEpi epi) {
  __shared__ pyn::SpmvSmem sm;
  for (int b = blockIdx.x; b < P.nblk; b += gridDim.x) pyn::spmv_stream_chunk(P, ind, val, x, b, sm, epi);
}

// bulk-copy variant (spmv.cuh): all chunks but the last through the TMA pipeline; the last chunk (whose aligned
// over-read could leave the arrays) goes through the register path in a one-CTA launch of spmv_tail_kernel
__global__ void __launch_bounds__(pyn::SB_THREADS, 4)
spmv_bulk_kernel(pyn::SpmvPlanView P, const int* __restrict__ ind, const double* __restrict__ val,
                 const double* __restrict__ x, ModeEpi epi) {
  extern __shared__ __align__(16) unsigned char sb_raw[];
  pyn::SbSmem& sm = *reinterpret_cast<pyn::SbSmem*>(sb_raw);
  pyn::spmv_bulk_loop(P, ind, val, x, blockIdx.x, gridDim.x, P.nbblk - 1, sm, epi);
}
__global__ void __launch_bounds__(pyn::SP_BLK)
spmv_tail_kernel(pyn::SpmvPlanView P, const int* __restrict__ ind, const double* __restrict__ val,
                 const double* __restrict__ x, ModeEpi epi) {
  __shared__ pyn::SpmvSmem sm;
  pyn::spmv_stream_rows(P, ind, val, x, P.brow[P.nbblk - 1], P.brow[P.nbblk], sm, epi);
}

__global__ void plan_p0_kernel(int nblk, const int64_t* __restrict__ ptr, const int* __restrict__ blk_row,
                               long long* __restrict__ blk_p0, int cap, int* __restrict__ too_big) {
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b <= nblk; b += gridDim.x * blockDim.x) {
    blk_p0[b] = ptr[blk_row[b]];
    if (b < nblk && ptr[blk_row[b + 1]] - ptr[blk_row[b]] > cap) atomicExch(too_big, 1);
  }
}

__global__ void plan_maxrow_kernel(int64_t nrows, const int64_t* __restrict__ ptr, int* __restrict__ out) {
  int m = 0;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t len = ptr[r + 1] - ptr[r];
    m = max(m, (int)min(len, (int64_t)0x7fffffff));
  }
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_down_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

// Chunks are cut by WEIGHT: a stored entry counts 1, a row counts `w`, so that a chunk holds at most `chunk` entries
// AND at most chunk / w rows -- a run of empty rows (boundary-only matrices, the restriction of one strip) is spread over
// many CTAs instead of landing in one.  blk_row[b] = first row whose weight offset ptr[r] + r w is >= b * chunk.
__global__ void plan_rows_kernel(int64_t nrows, const int64_t* __restrict__ ptr, int chunk, int w, int nblk,
                                 int* __restrict__ blk_row) {
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b <= nblk; b += gridDim.x * blockDim.x) {
    if (b == nblk) { blk_row[b] = (int)nrows; continue; }
    const int64_t key = (int64_t)b * chunk;
    int64_t lo = 0, hi = nrows;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (ptr[mid] + mid * w < key) lo = mid + 1; else hi = mid;
    }
    blk_row[b] = (int)lo;
  }
}

}  // namespace

extern "C" int pyn_spmv_plan_create(int64_t nrows, const int64_t* d_ptr, int64_t nnz, void* stream,
                                    pyn_spmv_plan** out) {
  PYN_REQUIRE(nrows > 0 && nrows < 2147483647LL && d_ptr && nnz >= 0 && out, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  pyn_spmv_plan* p = new pyn_spmv_plan();
  memset(p, 0, sizeof(*p));
  p->nrows = nrows; p->nnz = nnz; p->ptr = d_ptr;
  int* d_max = nullptr;
  int maxrow = 0;
  PYN_CHECK_CUDA(cudaMalloc(&d_max, sizeof(int)));
  PYN_CHECK_CUDA(cudaMemsetAsync(d_max, 0, sizeof(int), st));
  plan_maxrow_kernel<<<pyn::grid_for(nrows, 256), 256, 0, st>>>(nrows, d_ptr, d_max);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaMemcpyAsync(&maxrow, d_max, sizeof(int), cudaMemcpyDeviceToHost, st));
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  cudaFree(d_max);
  // one pass of SP_U loads per thread covers SP_U*SP_BLK products: aim chunks at that, leaving room for the last row
  const int pass = pyn::SP_U * pyn::SP_BLK;
  int chunk = pass - maxrow;
  if (chunk < pass / 2) chunk = pass / 2;
  p->chunk = chunk;
  const int roww = chunk / pyn::SP_RCAP > 1 ? chunk / pyn::SP_RCAP : 1;       // rows per chunk <= SP_RCAP
  const int64_t nb = (nnz + nrows * roww + chunk - 1) / chunk;
  PYN_REQUIRE(nb < 2147483647LL, "matrix too large for one SpMV plan");
  p->nblk = (int)nb;
  PYN_CHECK_CUDA(cudaMalloc(&p->blk_row, sizeof(int) * (nb + 1)));
  plan_rows_kernel<<<pyn::grid_for(nb + 1, 256), 256, 0, st>>>(nrows, d_ptr, chunk, roww, p->nblk, p->blk_row);
  PYN_CHECK_LAUNCH();
  {  // bulk-copy pipeline: its own finer chunk table, first non-zero of every chunk, whether every chunk fits a stage
    int bchunk = pyn::SB_CHUNK - maxrow;
    const bool fits = bchunk >= pyn::SB_CHUNK / 2;
    if (!fits) bchunk = pyn::SB_CHUNK / 2;
    const int broww = bchunk / pyn::SB_RCAP > 1 ? bchunk / pyn::SB_RCAP : 1;
    const int64_t nbb = (nnz + nrows * broww + bchunk - 1) / bchunk;
    PYN_REQUIRE(nbb < 2147483647LL, "matrix too large for one SpMV plan");
    p->nbblk = (int)nbb;
    PYN_CHECK_CUDA(cudaMalloc(&p->brow, sizeof(int) * (nbb + 1)));
    plan_rows_kernel<<<pyn::grid_for(nbb + 1, 256), 256, 0, st>>>(nrows, d_ptr, bchunk, broww, p->nbblk, p->brow);
    PYN_CHECK_LAUNCH();
    PYN_CHECK_CUDA(cudaMalloc(&p->bp0, sizeof(long long) * (nbb + 1)));
    int* d_big = nullptr;
    int big = 0;
    PYN_CHECK_CUDA(cudaMalloc(&d_big, sizeof(int)));
    PYN_CHECK_CUDA(cudaMemsetAsync(d_big, 0, sizeof(int), st));
    plan_p0_kernel<<<pyn::grid_for(nbb + 1, 256), 256, 0, st>>>(p->nbblk, d_ptr, p->brow, p->bp0, pyn::SB_CAP - 8, d_big);
    PYN_CHECK_LAUNCH();
    PYN_CHECK_CUDA(cudaMemcpyAsync(&big, d_big, sizeof(int), cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_big);
    // Measured on B200 at 1024^2 x ngl 5 (29.8 GB per product): register-staged kernel 6.23 ms (4.78 TB/s); this
    // pipeline 7.4 ms with 2048-entry chunks / 2 stages / 3 CTAs per SM, 8.9 ms with 3 stages / 2 CTAs, 9.2 ms with
    // 1024-entry chunks / 4 stages / 4 CTAs (profiles/README.md).  It therefore stays opt-in: PYN_SPMV_BULK=1.
    static const bool on = getenv("PYN_SPMV_BULK") ? atoi(getenv("PYN_SPMV_BULK")) != 0 : false;
    p->bulk_ok = fits && !big && on && p->nbblk >= 2 && (reinterpret_cast<uintptr_t>(d_ptr) & 15) == 0;
    PYN_CHECK_CUDA(cudaFuncSetAttribute(spmv_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(pyn::SbSmem)));
    int occb = 0;
    PYN_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occb, spmv_bulk_kernel, pyn::SB_THREADS, sizeof(pyn::SbSmem)));
    const int64_t capb = (int64_t)pyn::sm_count() * (occb > 0 ? occb : 1);
    p->bulk_grid = (int)(nbb < capb ? nbb : capb);
  }
  int occ = 0;  // persistent grid: exactly the CTAs that are resident at once
  PYN_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, spmv_stream_kernel, pyn::SP_BLK, 0));
  const int64_t cap = (int64_t)pyn::sm_count() * (occ > 0 ? occ : 1);
  p->grid = (int)(nb < cap ? nb : cap);
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  *out = p;
  return PYN_OK;
}

extern "C" int pyn_spmv_plan_destroy(pyn_spmv_plan* p) {
  if (!p) return PYN_OK;
  cudaFree(p->blk_row);
  cudaFree(p->brow);
  cudaFree(p->bp0);
  delete p;
  return PYN_OK;
}

extern "C" int pyn_spmv_planned(const pyn_spmv_plan* p, const int32_t* d_ind, const double* d_val, const double* d_x,
                                double* d_y, int mode, double p0, double p1, const double* d_aux, void* stream) {
  PYN_REQUIRE(p && d_ind && d_val && d_x && d_y, "bad arguments");
  PYN_REQUIRE(mode >= 0 && mode <= 3, "bad mode");
  PYN_REQUIRE(mode != PYN_SPMV_SRT_VTENSV || (d_aux != nullptr && p->nrows % 3 == 0),
              "SRT_VTENSV needs aux and 3 rows per node");
  ModeEpi epi{d_y, mode, p0, p1, d_aux};
  if (p->bulk_ok && ((reinterpret_cast<uintptr_t>(d_ind) | reinterpret_cast<uintptr_t>(d_val)) & 15) == 0) {
    spmv_bulk_kernel<<<p->bulk_grid, pyn::SB_THREADS, sizeof(pyn::SbSmem), pyn::as_stream(stream)>>>(pyn::view_of(p), d_ind,
                                                                                                    d_val, d_x, epi);
    pyn::count_launch();
    spmv_tail_kernel<<<1, pyn::SP_BLK, 0, pyn::as_stream(stream)>>>(pyn::view_of(p), d_ind, d_val, d_x, epi);
  } else
    spmv_stream_kernel<<<p->grid, pyn::SP_BLK, 0, pyn::as_stream(stream)>>>(pyn::view_of(p), d_ind, d_val, d_x, epi);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

extern "C" int pyn_spmv(int64_t nrows, const int64_t* d_ptr, const int32_t* d_ind, const double* d_val, int64_t nnz,
                        const double* d_x, double* d_y, int mode, double p0, double p1, const double* d_aux,
                        void* stream) {
  PYN_REQUIRE(nrows > 0 && d_ptr && d_ind && d_val && d_x && d_y, "bad arguments");
  PYN_REQUIRE(mode >= 0 && mode <= 3, "bad mode");
  PYN_REQUIRE(mode != PYN_SPMV_SRT_VTENSV || (d_aux != nullptr && nrows % 3 == 0), "SRT_VTENSV needs aux and 3 rows per node");
  cudaStream_t st = pyn::as_stream(stream);
  switch (pyn::pick_lpr(nrows, nnz)) {
    case 4: return launch<4>(nrows, d_ptr, d_ind, d_val, d_x, d_y, mode, p0, p1, d_aux, st);
    case 8: return launch<8>(nrows, d_ptr, d_ind, d_val, d_x, d_y, mode, p0, p1, d_aux, st);
    case 16: return launch<16>(nrows, d_ptr, d_ind, d_val, d_x, d_y, mode, p0, p1, d_aux, st);
    default: return launch<32>(nrows, d_ptr, d_ind, d_val, d_x, d_y, mode, p0, p1, d_aux, st);
  }
}
