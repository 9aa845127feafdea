// Element operator kernels (see tmat.cuh): y = A x with fused epilogues, same modes as pyn_spmv.
#include "tmat.cuh"

namespace {

using pyn::EO_Q;
using pyn::EO_R;
using pyn::EO_TE;
using pyn::EO_THREADS;

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

// E (row major [NR][NC]) -> sign * E^T, rows padded to NRP with zeros
__global__ void eop_transpose_kernel(int NR, int NC, int NRP, double sign, const double* __restrict__ E,
                                     double* __restrict__ Et) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < NC * NRP; i += gridDim.x * blockDim.x) {
    const int k = i / NRP, r = i - k * NRP;
    Et[i] = (r < NR) ? sign * E[r * NC + k] : 0.0;
  }
}

// Phase 1.  Persistent CTAs; E^T stays in shared memory for the life of the CTA.  Per tile of 128 elements:
//   gather   lane <-> element (conflict-free smem stores), warps split the local nodes; x through the read-only path
//   compute  work item = 8 rows x (4 elements per lane); per k: 2 LDS.128 of X + 4 broadcast LDS.128 of E^T feed 32 DFMA
//   store    each thread writes its 8 consecutive rows of 4 elements (64 B runs of the element-major Ye)
template <int CD>
__global__ void __launch_bounds__(EO_THREADS)
eop_apply_kernel(pyn::TmatView m, const double* __restrict__ x, const int* __restrict__ done) {
  extern __shared__ __align__(16) double sm[];
  if (done && *done) return;
  const int NR = m.NR, NC = m.NC, NRP = m.NRP, nn = m.nn;
  double* Et = sm;                  // [NC][NRP]
  double* Xs = sm + NC * NRP;       // [NC][EO_TE]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < NC * NRP; i += EO_THREADS) Et[i] = m.Et[i];
  const int64_t ntiles = (m.ne + EO_TE - 1) / EO_TE;
  const int nitems = NRP / EO_R;
  const bool vec_store = (NR & 1) == 0;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t e0 = tile * EO_TE;
    __syncthreads();   // the previous tile's products are done (and E^T is in place)
    for (int j = warp; j < nn; j += EO_THREADS / 32) {
#pragma unroll
      for (int q = 0; q < EO_TE / 32; ++q) {
        const int el = q * 32 + lane;
        const int64_t e = e0 + el;
        double v[CD];
#pragma unroll
        for (int b = 0; b < CD; ++b) v[b] = 0.0;
        if (e < m.ne) {
          const int64_t node = m.e2n[e * nn + j];
          if (CD == 2) {
            const double2 t = __ldg(reinterpret_cast<const double2*>(x + node * 2));
            v[0] = t.x;
            v[CD - 1] = t.y;
          } else {
#pragma unroll
            for (int b = 0; b < CD; ++b) v[b] = __ldg(x + node * CD + b);
          }
        }
#pragma unroll
        for (int b = 0; b < CD; ++b) Xs[(j * CD + b) * EO_TE + el] = v[b];
      }
    }
    __syncthreads();
    for (int item = warp; item < nitems; item += EO_THREADS / 32) {
      const int i0 = item * EO_R;
      double acc[EO_R][EO_Q];
#pragma unroll
      for (int r = 0; r < EO_R; ++r)
#pragma unroll
        for (int q = 0; q < EO_Q; ++q) acc[r][q] = 0.0;
      const double* xp = Xs + EO_Q * lane;
      const double* ep = Et + i0;
#pragma unroll 2
      for (int k = 0; k < NC; ++k) {
        const double2 xa = *reinterpret_cast<const double2*>(xp + k * EO_TE);
        const double2 xb = *reinterpret_cast<const double2*>(xp + k * EO_TE + 2);
        const double xv[EO_Q] = {xa.x, xa.y, xb.x, xb.y};
        double ev[EO_R];
#pragma unroll
        for (int r = 0; r < EO_R; r += 2) {
          const double2 t = *reinterpret_cast<const double2*>(ep + k * NRP + r);
          ev[r] = t.x;
          ev[r + 1] = t.y;
        }
#pragma unroll
        for (int r = 0; r < EO_R; ++r)
#pragma unroll
          for (int q = 0; q < EO_Q; ++q) acc[r][q] = fma(ev[r], xv[q], acc[r][q]);
      }
#pragma unroll
      for (int q = 0; q < EO_Q; ++q) {
        const int64_t e = e0 + EO_Q * lane + q;
        if (e >= m.ne) continue;
        double* yp = m.Ye + e * NR + i0;
        if (vec_store) {
#pragma unroll
          for (int r = 0; r < EO_R; r += 2)
            if (i0 + r < NR) *reinterpret_cast<double2*>(yp + r) = make_double2(acc[r][q], acc[r + 1][q]);
        } else {
#pragma unroll
          for (int r = 0; r < EO_R; ++r)
            if (i0 + r < NR) yp[r] = acc[r][q];
        }
      }
    }
  }
}

// Phase 2: one thread per un-flagged node
template <int RD>
__global__ void __launch_bounds__(256) eop_gather_kernel(pyn::TmatView m, ModeEpi epi) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < m.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    if (m.flags[r]) continue;
    double acc[RD];
    pyn::eop_gather_node<RD>(m, r, acc);
#pragma unroll
    for (int a = 0; a < RD; ++a) epi(r * RD + a, acc[a]);
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

size_t eop_smem(int NC, int NRP) { return sizeof(double) * ((size_t)NC * NRP + (size_t)NC * EO_TE); }
constexpr size_t EO_SMEM_MAX = 227 * 1024;

}  // namespace

namespace pyn {

int tmat_apply(const pyn_tmat* m, const double* d_x, const int* d_done, cudaStream_t st) {
  const TmatView v = view_of(m);
  if (m->cdof == 1) eop_apply_kernel<1><<<m->grid_apply, EO_THREADS, m->smem, st>>>(v, d_x, d_done);
  else if (m->cdof == 2) eop_apply_kernel<2><<<m->grid_apply, EO_THREADS, m->smem, st>>>(v, d_x, d_done);
  else eop_apply_kernel<3><<<m->grid_apply, EO_THREADS, m->smem, st>>>(v, d_x, d_done);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

// called by pyn_tmat_create (assemble.cu, which owns the descriptor logic): element matrix, scratch, launch shapes
int tmat_init(pyn_tmat* m, const double* d_E, double sign, cudaStream_t st) {
  const pyn_topo* t = m->topo;
  m->NR = m->rdof * t->nn; m->NC = m->cdof * t->nn;
  m->NRP = (m->NR + EO_R - 1) / EO_R * EO_R;
  m->smem = eop_smem(m->NC, m->NRP);
  PYN_REQUIRE(m->smem <= EO_SMEM_MAX, "element operator: the element matrix does not fit in shared memory");
  PYN_CHECK_CUDA(cudaMalloc(&m->Et, sizeof(double) * (size_t)m->NC * m->NRP));
  PYN_CHECK_CUDA(cudaMalloc(&m->Ye, sizeof(double) * (size_t)t->ne * m->NR));
  eop_transpose_kernel<<<(m->NC * m->NRP + 255) / 256, 256, 0, st>>>(m->NR, m->NC, m->NRP, sign, d_E, m->Et);
  PYN_CHECK_LAUNCH();
  int occ = 0;
#define PYN_EO_ATTR(CD)                                                                                            \
  PYN_CHECK_CUDA(cudaFuncSetAttribute(eop_apply_kernel<CD>, cudaFuncAttributeMaxDynamicSharedMemorySize,           \
                                      (int)EO_SMEM_MAX));                                                          \
  PYN_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, eop_apply_kernel<CD>, EO_THREADS, m->smem));
  if (m->cdof == 1) { PYN_EO_ATTR(1) } else if (m->cdof == 2) { PYN_EO_ATTR(2) } else { PYN_EO_ATTR(3) }
#undef PYN_EO_ATTR
  const int64_t ntiles = (t->ne + EO_TE - 1) / EO_TE;
  const int64_t cap_a = (int64_t)sm_count() * (occ > 0 ? occ : 1);
  m->grid_apply = (int)(ntiles < cap_a ? ntiles : cap_a);
  m->grid_int = grid_for(t->nnodes, 256, 8);
  const int64_t need_b = (m->nlist * m->rdof * 32 + 255) / 256, cap_b = (int64_t)sm_count() * 2;
  m->grid_bnd = (int)(need_b < cap_b ? (need_b > 0 ? need_b : 1) : cap_b);
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  return PYN_OK;
}

}  // namespace pyn

extern "C" {

int pyn_tmat_supported(int nn, int rdof, int cdof) {
  if (nn < 1 || rdof < 1 || rdof > 3 || cdof < 1 || cdof > 3) return 0;
  const int NR = rdof * nn, NC = cdof * nn, NRP = (NR + EO_R - 1) / EO_R * EO_R;
  return eop_smem(NC, NRP) <= EO_SMEM_MAX ? 1 : 0;
}

int pyn_tmat_destroy(pyn_tmat* m) {
  if (!m) return PYN_OK;
  cudaFree(m->Et);
  cudaFree(m->Ye);
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
  int rc = pyn::tmat_apply(m, d_x, nullptr, st);
  if (rc) return rc;
  if (m->nlist > 0) {
    tmat_boundary_kernel<<<m->grid_bnd, 256, 0, st>>>(v, m->rdof, d_x, epi);
    PYN_CHECK_LAUNCH();
  }
  if (m->rdof == 1) eop_gather_kernel<1><<<m->grid_int, 256, 0, st>>>(v, epi);
  else if (m->rdof == 2) eop_gather_kernel<2><<<m->grid_int, 256, 0, st>>>(v, epi);
  else eop_gather_kernel<3><<<m->grid_int, 256, 0, st>>>(v, epi);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

/* algorithmic HBM bytes of one product (tmat.cuh): x once, connectivity, Ye written and read once, the incidence
 * lists, y once; plus the CSR rows of the flagged nodes at 12 B per entry (not counted: O(boundary)) */
int64_t pyn_tmat_bytes(const pyn_tmat* m) {
  if (!m) return 0;
  const pyn_topo* t = m->topo;
  const int64_t inc = t->ne * (int64_t)t->nn;
  return 8 * t->nnodes * m->cdof + 4 * inc + 16 * t->ne * (int64_t)m->NR + 8 * (t->nnodes + 1) + 6 * inc +
         8 * t->nnodes * m->rdof;
}
}
