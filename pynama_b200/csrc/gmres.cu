// Restarted GMRES on device with PETSc's KSPGMRES semantics (-ksp_type gmres; KspSolver, src/solver/ksp_solver.py:9-19):
// left preconditioning (Jacobi or none), zero (or given: nonzero_guess) initial guess, restart 30 by default, CLASSICAL Gram-Schmidt without
// refinement (all projections from the same vector), residual norm from the Givens recurrence, KSPConvergedDefault on the
// preconditioned residual; at the top of every cycle the true preconditioned residual is formed and tested.
//
// The Krylov basis lives in HBM ((restart+1) vectors); every long-vector operation is a kernel (CSR-stream SpMV with the
// preconditioner in its epilogue, blocked multi-dot, fused projection + norm, scaling); only the (restart+1)^2 Hessenberg
// matrix and its Givens rotations are on the host, which costs one small device->host copy per iteration.
#include <math.h>

#include <vector>

#include "ksp.cuh"

namespace {
constexpr int BLK = 256;
constexpr int MD = 8;  // basis vectors handled per multi-dot / multi-axpy launch

struct VecPack {
  const double* v[MD];
  double c[MD];
  int m;
};

// w = M^-1 (A v)           (resid = false)
// w = M^-1 (b - A v)       (resid = true)
struct PcEpi {
  double* w;
  const double* dinv;
  const double* b;
  const uint8_t* owned;
  __device__ __forceinline__ void operator()(int64_t row, double s) const {
    double r = b ? (b[row] - s) : s;
    if (owned && !owned[row]) r = 0.0;
    w[row] = dinv[row] * r;
  }
};
__global__ void __launch_bounds__(pyn::SP_BLK, 6)
gmres_spmv_kernel(pyn::SpmvPlanView P, const int* __restrict__ ind, const double* __restrict__ val,
                  const double* __restrict__ x, PcEpi epi) {
  __shared__ pyn::SpmvSmem sm;
  for (int b = blockIdx.x; b < P.nblk; b += gridDim.x) pyn::spmv_stream_chunk(P, ind, val, x, b, sm, epi);
}

// the same epilogue after a product through the element operator (pyn_tmat_mult wrote w = A v)
__global__ void gmres_pc_kernel(int64_t n, PcEpi epi) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) epi(i, epi.w[i]);
}

// r = M^-1 b (zero initial guess), x = 0
__global__ void gmres_init_kernel(int64_t n, const double* __restrict__ b, const double* __restrict__ dinv,
                                  const uint8_t* __restrict__ owned, double* __restrict__ r, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    r[i] = (owned && !owned[i]) ? 0.0 : dinv[i] * b[i];
    x[i] = 0.0;
  }
}

// partial[blk][j] = sum_i v_j[i] * w[i]   for the pack's m vectors  (w.w when v_j == NULL)
// `owned` (multi-GPU): ghost entries are another rank's and are left out of the sums
__global__ void gmres_mdot_kernel(int64_t n, const __grid_constant__ VecPack pk, const double* __restrict__ w,
                                  const uint8_t* __restrict__ owned, double* __restrict__ part) {
  __shared__ double sm[BLK / 32];
  double acc[MD];
#pragma unroll
  for (int j = 0; j < MD; ++j) acc[j] = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    if (owned && !owned[i]) continue;
    const double wi = w[i];
#pragma unroll
    for (int j = 0; j < MD; ++j)
      if (j < pk.m) acc[j] += (pk.v[j] ? pk.v[j][i] : wi) * wi;
  }
#pragma unroll
  for (int j = 0; j < MD; ++j) {
    if (j < pk.m) {
      const double s = pyn::block_sum<BLK>(acc[j], sm);
      if (threadIdx.x == 0) part[(int64_t)blockIdx.x * MD + j] = s;
    }
  }
}
// out[j] = sum over blocks of part[blk][j]   (fixed order: deterministic)
__global__ void gmres_mdot_final_kernel(int nblk, int m, const double* __restrict__ part, double* __restrict__ out) {
  const int j = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (j >= m) return;
  double s = 0.0;
  for (int b = lane; b < nblk; b += 32) s += part[(int64_t)b * MD + j];
  s = pyn::warp_sum(s);
  if (lane == 0) out[j] = s;
}

// w -= sum_j h[j] v_j  with the coefficients read from device memory (no host round trip inside Gram-Schmidt)
__global__ void gmres_project_kernel(int64_t n, const __grid_constant__ VecPack pk, const double* __restrict__ h,
                                     double* __restrict__ w) {
  double c[MD];
#pragma unroll
  for (int j = 0; j < MD; ++j) c[j] = (j < pk.m) ? h[j] : 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    double s = w[i];
#pragma unroll
    for (int j = 0; j < MD; ++j)
      if (j < pk.m) s -= c[j] * pk.v[j][i];
    w[i] = s;
  }
}
// v = w / sqrt(nrm2[0])
__global__ void gmres_scale_kernel(int64_t n, const double* __restrict__ w, const double* __restrict__ nrm2,
                                   double* __restrict__ v) {
  const double t = sqrt(nrm2[0]);
  const double inv = t > 0.0 ? 1.0 / t : 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) v[i] = w[i] * inv;
}
// x += sum_j c_j v_j  (coefficients from the host: the solution of the small triangular system)
__global__ void gmres_axpy_kernel(int64_t n, const __grid_constant__ VecPack pk, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    double s = x[i];
#pragma unroll
    for (int j = 0; j < MD; ++j)
      if (j < pk.m) s += pk.c[j] * pk.v[j][i];
    x[i] = s;
  }
}
__global__ void gmres_copy_kernel(int64_t n, const double* __restrict__ src, double* __restrict__ dst) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) dst[i] = src[i];
}
}  // namespace

struct pyn_gmres {
  int m;           // restart
  double* V;       // (m+1) basis vectors
  double* w;       // work vector
  double* x;       // iterate
  double* part;    // multi-dot partials [vgrid][MD]
  double* dots;    // device scalars: h column (m+2)
  double* h_dots;  // pinned mirror
};

namespace pyn {

void gmres_free(pyn_gmres* g) {
  if (!g) return;
  cudaFree(g->V);
  cudaFree(g->part);
  cudaFree(g->dots);
  cudaFreeHost(g->h_dots);
  delete g;
}

#define GM_LAUNCH_CHECK()                                                                       \
  do {                                                                                          \
    ++launches;                                                                                 \
    cudaError_t _e = cudaGetLastError();                                                        \
    if (_e != cudaSuccess) {                                                                    \
      pyn::set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return PYN_ERR_CUDA;                                                                      \
    }                                                                                           \
  } while (0)

int gmres_solve(pyn_ksp* k, const double* d_b, double* d_x, pyn_ksp_result* res, cudaStream_t st) {
  const int64_t n = k->n;
  const int m = k->opts.restart > 0 ? k->opts.restart : 30;
  const bool dist = k->comm && k->comm->world > 1;
  const uint8_t* owned = dist ? k->owned : nullptr;
  int64_t launches = 0;
  if (!k->gm || k->gm->m != m) {
    gmres_free(k->gm);
    k->gm = nullptr;
    pyn_gmres* g = new pyn_gmres();
    memset(g, 0, sizeof(*g));
    g->m = m;
    PYN_CHECK_CUDA(cudaMalloc(&g->V, sizeof(double) * (size_t)n * (m + 3)));
    g->w = g->V + (size_t)n * (m + 1);
    g->x = g->w + n;
    PYN_CHECK_CUDA(cudaMalloc(&g->part, sizeof(double) * (size_t)k->vgrid * MD));
    PYN_CHECK_CUDA(cudaMalloc(&g->dots, sizeof(double) * (m + 4)));
    PYN_CHECK_CUDA(cudaMallocHost(&g->h_dots, sizeof(double) * (m + 4)));
    k->gm = g;
  }
  pyn_gmres* g = k->gm;
  const double* dinv = k->B.dinv;
  const pyn::SpmvPlanView P = pyn::view_of(k->plan);
  auto Vj = [&](int j) { return g->V + (size_t)n * j; };
  int rc;

  // dots[0..cnt) = <v_j, w> for j < cnt (v == NULL: <w, w>), all-reduced over the ranks; optionally copied to the host
  auto multi_dot = [&](int first, int cnt, const double* w, bool self, double* d_out) -> int {
    for (int j0 = 0; j0 < cnt; j0 += MD) {
      VecPack pk;
      pk.m = (cnt - j0 < MD) ? cnt - j0 : MD;
      for (int j = 0; j < MD; ++j) pk.v[j] = (j < pk.m && !self) ? Vj(first + j0 + j) : nullptr;
      gmres_mdot_kernel<<<k->vgrid, BLK, 0, st>>>(n, pk, w, owned, g->part);
      GM_LAUNCH_CHECK();
      gmres_mdot_final_kernel<<<1, 32 * MD, 0, st>>>(k->vgrid, pk.m, g->part, d_out + j0);
      GM_LAUNCH_CHECK();
    }
    if (dist) {
      int rc = pyn::comm_allreduce_sum(k->comm, d_out, cnt, st);
      if (rc) return rc;
    }
    return PYN_OK;
  };
  auto fetch = [&](int cnt) -> int {
    PYN_CHECK_CUDA(cudaMemcpyAsync(g->h_dots, g->dots, sizeof(double) * cnt, cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaStreamSynchronize(st));
    return PYN_OK;
  };

  // epi.w = M^-1 (A v)  or  M^-1 (b - A v): CSR stream with the epilogue fused, or element operator + one vector pass
  auto product = [&](const double* v, const PcEpi& epi) -> int {
    if (dist && (rc = pyn::halo_exchange(k->halo, const_cast<double*>(v), k->halo_comp, st))) return rc;
    if (k->tmat) {
      if ((rc = pyn_tmat_mult(k->tmat, v, epi.w, PYN_SPMV_SET, 0.0, 0.0, nullptr, st))) return rc;
      gmres_pc_kernel<<<k->vgrid, BLK, 0, st>>>(n, epi);
      launches += 3;
    } else {
      gmres_spmv_kernel<<<k->sgrid, pyn::SP_BLK, 0, st>>>(P, k->ind, k->val, v, epi);
    }
    GM_LAUNCH_CHECK();
    return PYN_OK;
  };

  const double rtol = k->opts.rtol, atol = k->opts.atol, dtol = k->opts.dtol;
  const int max_it = k->opts.max_it;
  const bool guess = k->opts.nonzero_guess != 0;
  double bnorm = 0.0;
  std::vector<double> H((size_t)(m + 1) * m, 0.0), cs(m, 0.0), sn(m, 0.0), gg(m + 1, 0.0), y(m, 0.0);
  int its = 0, reason = 0;
  double rnorm0 = 0.0, rnorm = 0.0, ttol = 0.0;
  bool first = true;

  while (!reason) {
    // ---- residual of the cycle: V0 = M^-1 (b - A x) ------------------------------------------------------------
    if (first) {
      gmres_init_kernel<<<k->vgrid, BLK, 0, st>>>(n, d_b, dinv, owned, Vj(0), g->x);
      GM_LAUNCH_CHECK();
    }
    if (first && guess) {   // -ksp_initial_guess_nonzero: |M^-1 b| is the yardstick of rtol, x0 = contents of d_x
      if ((rc = multi_dot(0, 1, Vj(0), true, g->dots))) return rc;
      if ((rc = fetch(1))) return rc;
      bnorm = sqrt(g->h_dots[0]);
      gmres_copy_kernel<<<k->vgrid, BLK, 0, st>>>(n, d_x, g->x);
      GM_LAUNCH_CHECK();
    }
    if (!first || guess) {
      PcEpi epi{Vj(0), dinv, d_b, owned};
      if ((rc = product(g->x, epi))) return rc;
    }
    if ((rc = multi_dot(0, 1, Vj(0), true, g->dots))) return rc;
    if ((rc = fetch(1))) return rc;
    const double beta = sqrt(g->h_dots[0]);
    rnorm = beta;
    if (first) {
      rnorm0 = (guess && bnorm > 0.0) ? bnorm : beta;
      ttol = fmax(rtol * rnorm0, atol);
      first = false;
    }
    if (rnorm <= ttol) { reason = (rnorm <= atol) ? 3 : 2; break; }   // KSPConvergedDefault on the true residual
    if (rnorm >= dtol * rnorm0) { reason = -4; break; }
    if (its >= max_it) { reason = -3; break; }
    gmres_scale_kernel<<<k->vgrid, BLK, 0, st>>>(n, Vj(0), g->dots, Vj(0));
    GM_LAUNCH_CHECK();
    std::fill(gg.begin(), gg.end(), 0.0);
    gg[0] = beta;
    int j = 0;
    for (; j < m; ++j) {
      // ---- w = M^-1 A v_j -----------------------------------------------------------------------------------
      PcEpi epi{g->w, dinv, nullptr, owned};
      if ((rc = product(Vj(j), epi))) return rc;
      // ---- classical Gram-Schmidt: h_i = <v_i, w> for all i <= j from the same w, then one projection ----------
      if ((rc = multi_dot(0, j + 1, g->w, false, g->dots))) return rc;
      for (int j0 = 0; j0 <= j; j0 += MD) {
        VecPack pk;
        pk.m = (j + 1 - j0 < MD) ? j + 1 - j0 : MD;
        for (int q = 0; q < MD; ++q) pk.v[q] = q < pk.m ? Vj(j0 + q) : nullptr;
        gmres_project_kernel<<<k->vgrid, BLK, 0, st>>>(n, pk, g->dots + j0, g->w);
        GM_LAUNCH_CHECK();
      }
      if ((rc = multi_dot(0, 1, g->w, true, g->dots + j + 1))) return rc;
      gmres_scale_kernel<<<k->vgrid, BLK, 0, st>>>(n, g->w, g->dots + j + 1, Vj(j + 1));
      GM_LAUNCH_CHECK();
      if ((rc = fetch(j + 2))) return rc;
      // ---- Hessenberg column, Givens rotations (host) -----------------------------------------------------------
      double* hc = &H[(size_t)j * (m + 1)];
      for (int i = 0; i <= j; ++i) hc[i] = g->h_dots[i];
      const double hnext = sqrt(g->h_dots[j + 1]);
      hc[j + 1] = hnext;
      for (int i = 0; i < j; ++i) {
        const double t = cs[i] * hc[i] + sn[i] * hc[i + 1];
        hc[i + 1] = -sn[i] * hc[i] + cs[i] * hc[i + 1];
        hc[i] = t;
      }
      const double den = sqrt(hc[j] * hc[j] + hc[j + 1] * hc[j + 1]);
      if (den == 0.0) { reason = -5; ++its; break; }
      cs[j] = hc[j] / den;
      sn[j] = hc[j + 1] / den;
      hc[j] = den;
      hc[j + 1] = 0.0;
      gg[j + 1] = -sn[j] * gg[j];
      gg[j] = cs[j] * gg[j];
      rnorm = fabs(gg[j + 1]);
      ++its;
      if (rnorm <= ttol) reason = (rnorm <= atol) ? 3 : 2;
      else if (hnext == 0.0) reason = 2;                     // happy breakdown
      else if (rnorm >= dtol * rnorm0) reason = -4;
      else if (its >= max_it) reason = -3;
      if (reason) { ++j; break; }
    }
    // ---- x += V y with  R y = g  (back substitution on the host) --------------------------------------------------
    const int kdim = j;   // number of basis vectors used in this cycle
    for (int i = kdim - 1; i >= 0; --i) {
      double s = gg[i];
      for (int q = i + 1; q < kdim; ++q) s -= H[(size_t)q * (m + 1) + i] * y[q];
      y[i] = s / H[(size_t)i * (m + 1) + i];
    }
    for (int j0 = 0; j0 < kdim; j0 += MD) {
      VecPack pk;
      pk.m = (kdim - j0 < MD) ? kdim - j0 : MD;
      for (int q = 0; q < MD; ++q) {
        pk.v[q] = q < pk.m ? Vj(j0 + q) : nullptr;
        pk.c[q] = q < pk.m ? y[j0 + q] : 0.0;
      }
      gmres_axpy_kernel<<<k->vgrid, BLK, 0, st>>>(n, pk, g->x);
      GM_LAUNCH_CHECK();
    }
  }
  gmres_copy_kernel<<<k->vgrid, BLK, 0, st>>>(n, g->x, d_x);
  GM_LAUNCH_CHECK();
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  pyn::count_launch((int)launches);
  if (res) {
    res->its = its;
    res->reason = reason;
    res->rnorm0 = rnorm0;
    res->rnorm = rnorm;
  }
  k->last_spmv = its;
  k->last_launches = launches;
  return PYN_OK;
}

}  // namespace pyn
