// Template operator kernels (see tmat.cuh): y = A x with fused epilogues, same modes as pyn_spmv.
#include "tmat.cuh"

namespace {

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

template <int RD, int CD>
__global__ void __launch_bounds__(256)
tmat_interior_kernel(pyn::TmatView m, const double* __restrict__ x, ModeEpi epi) {
  const int lane = threadIdx.x & 31;
  const int64_t w0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = w0; r < m.nnodes; r += nw) {
    if (m.flags[r]) continue;
    double acc[RD];
    pyn::tmat_node<RD, CD>(m, r, lane, x, acc);
    if (lane == 0) {
#pragma unroll
      for (int a = 0; a < RD; ++a) epi(r * RD + a, acc[a]);
    }
  }
}

// rows of the flagged nodes from the assembled CSR: one warp per row
__global__ void __launch_bounds__(256)
tmat_boundary_kernel(pyn::TmatView m, int rd, const double* __restrict__ x, ModeEpi epi) {
  const int lane = threadIdx.x & 31;
  const int64_t w0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const int64_t nrows = m.nlist * rd;
  for (int64_t i = w0; i < nrows; i += nw) {
    const int64_t row = (int64_t)m.list[i / rd] * rd + (i % rd);
    const double s = pyn::spmv_row<32>(m.ptr, m.ind, m.val, x, row, true, lane);
    if (lane == 0) epi(row, s);
  }
}

}  // namespace

extern "C" {

int pyn_tmat_destroy(pyn_tmat* m) {
  if (!m) return PYN_OK;
  cudaFree(m->T);
  delete m;
  return PYN_OK;
}

int pyn_tmat_mult(const pyn_tmat* m, const double* d_x, double* d_y, int mode, double p0, double p1, const double* d_aux,
                  void* stream) {
  PYN_REQUIRE(m && d_x && d_y, "bad arguments");
  PYN_REQUIRE(mode >= 0 && mode <= 3, "bad mode");
  PYN_REQUIRE(mode != PYN_SPMV_SRT_VTENSV || (d_aux != nullptr && m->rdof == 3), "SRT_VTENSV needs aux and 3 rows per node");
  cudaStream_t st = pyn::as_stream(stream);
  const pyn::TmatView v = pyn::view_of(m);
  ModeEpi epi{d_y, mode, p0, p1, d_aux};
  if (m->nlist > 0) {
    tmat_boundary_kernel<<<m->grid_bnd, 256, 0, st>>>(v, m->rdof, d_x, epi);
    PYN_CHECK_LAUNCH();
  }
  bool done = false;
#define PYN_T_CASE(R, C)                                                        \
  if (!done && m->rdof == R && m->cdof == C) {                                  \
    tmat_interior_kernel<R, C><<<m->grid_int, 256, 0, st>>>(v, d_x, epi);       \
    done = true;                                                                \
  }
  PYN_T_CASE(2, 2) PYN_T_CASE(2, 1) PYN_T_CASE(1, 2) PYN_T_CASE(3, 2) PYN_T_CASE(2, 3)
  PYN_T_CASE(1, 1) PYN_T_CASE(3, 3) PYN_T_CASE(1, 3) PYN_T_CASE(3, 1)
#undef PYN_T_CASE
  PYN_REQUIRE(done, "template operator: unsupported block shape");
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

/* algorithmic HBM bytes of one product: 4 B per adjacency slot of the interior nodes (+ 8 B row pointer per node),
 * the CSR rows of the flagged nodes at 12 B per entry, the vectors once */
int64_t pyn_tmat_bytes(const pyn_tmat* m) {
  if (!m) return 0;
  const pyn_topo* t = m->topo;
  return 4 * t->adj_nnz + 8 * (t->nnodes + 1) + 8 * t->nnodes * (m->rdof + m->cdof);
}
}
