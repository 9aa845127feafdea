// y = A x with fused epilogues (see include/pynama_b200.h: pyn_spmv).
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

}  // namespace

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
