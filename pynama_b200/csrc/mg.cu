// Geometric multigrid V-cycle as a CG preconditioner for the velocity-from-vorticity operator K.
//
// The reference hides the cost of the solve behind a factor-once sparse LU (src/solver/ksp_solver.py:13-16).  On the
// device the solve is CG; with Jacobi its iteration count grows like 1/h (10 453 iterations at 1024^2 x ngl 5, rtol
// 1e-5).  The V-cycle makes it (almost) mesh independent:
//
//   levels   the fine space (nx x ny elements of order ngl), then the same mesh at lower orders (p-coarsening, e.g.
//            5 -> 3 -> 2), then bilinear elements on meshes coarsened by 2 in each direction (h-coarsening); every
//            level's operator is the REDISCRETISED K of that space -- assembled by the same integration and assembly
//            kernels and applied through the same element operator (tmat) or CSR stream as the fine one;
//   smoother Chebyshev polynomial of degree `degree` in D^-1 A on [lmax / ratio, lmax] (first-kind recurrence, the
//            one of PETSc's KSPCHEBYSHEV), the same polynomial before and after the coarse correction, so the cycle
//            is a symmetric positive definite operator and CG applies;
//   transfer nodal interpolation P of the coarse basis at the fine nodes, restriction R = P^T, both stored as CSR on
//            the unknowns (prescribed dofs carry no entries) and applied with the CSR-stream kernel;
//   coarsest dense inverse applied as a matrix-vector product.
//
// Prescribed (Dirichlet / wall) dofs are identity rows of K, decoupled from the unknowns (their couplings live in Krhs):
// every vector of the cycle is zero there (dinv = 0), the caller preconditions those rows with Jacobi.
// Nothing here synchronises or allocates after the first call: the whole cycle is captured inside the CG graph.
#include "mg.cuh"

namespace {
constexpr int BLK = 256;

// d = s * dinv * b ; x (=|+=) d ; optionally r = b (start of a Chebyshev sequence)
__global__ void mg_start_kernel(int64_t n, const double* __restrict__ b, const double* __restrict__ dinv, double s,
                                double* __restrict__ x, double* r, double* __restrict__ d, int zero_x, int copy_r) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const double bi = b[i];
    const double di = s * dinv[i] * bi;
    d[i] = di;
    x[i] = zero_x ? di : x[i] + di;
    if (copy_r) r[i] = bi;
  }
}
// r -= w ; d = c1 d + c2 dinv r ; x += d
__global__ void mg_step_kernel(int64_t n, const double* __restrict__ w, const double* __restrict__ dinv, double c1,
                               double c2, double* __restrict__ r, double* __restrict__ d, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const double ri = r[i] - w[i];
    const double di = c1 * d[i] + c2 * dinv[i] * ri;
    r[i] = ri;
    d[i] = di;
    x[i] += di;
  }
}
// r -= w ; optionally x += d (coarse-grid correction)
__global__ void mg_resid_kernel(int64_t n, const double* __restrict__ w, double* __restrict__ r,
                                const double* __restrict__ d, double* __restrict__ x) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    r[i] -= w[i];
    if (d) x[i] += d[i];
  }
}
// x = Ainv b, one warp per row (coarsest level)
__global__ void mg_dense_kernel(int n, const double* __restrict__ Ainv, const double* __restrict__ b,
                                double* __restrict__ x) {
  const int lane = threadIdx.x & 31;
  const int row = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5);
  if (row >= n) return;
  const double* a = Ainv + (int64_t)row * n;
  double s = 0.0;
  for (int j = lane; j < n; j += 32) s += a[j] * b[j];
  s = pyn::warp_sum(s);
  if (lane == 0) x[row] = s;
}

// y = A_l x.  On a partitioned fine level the ghost entries of x are refreshed from their owners first (x is one of
// the level's work vectors: only its ghost entries change).
int apply_op(const pyn_mg* g, const MgLevel& L, double* x, double* y, cudaStream_t st) {
  (void)g;
  if (L.halo) {
    int rc = pyn::halo_exchange(L.halo, x, L.halo_comp, st);
    if (rc) return rc;
  }
  if (L.tmat) return pyn_tmat_mult(L.tmat, x, y, PYN_SPMV_SET, 0.0, 0.0, nullptr, st);
  return pyn_spmv_planned(L.plan, L.ind, L.val, x, y, PYN_SPMV_SET, 0.0, 0.0, nullptr, st);
}

// `deg` Chebyshev steps on A x = b.  fresh: x = 0 and r = b on entry (b may alias r); otherwise r = b - A x is current.
// want_resid: leave r = b - A x up to date on exit (one more operator application).
int smooth(const pyn_mg* g, MgLevel& L, const double* b, double* x, bool fresh, bool want_resid, cudaStream_t st) {
  const double lmax = L.lmax, lmin = L.lmax / g->ratio;
  const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma = theta / delta;
  double rho = 1.0 / sigma;
  mg_start_kernel<<<L.vgrid, BLK, 0, st>>>(L.n, fresh ? b : L.r, L.dinv, 1.0 / theta, x, L.r, L.d, fresh ? 1 : 0,
                                           (fresh && b != L.r) ? 1 : 0);
  PYN_CHECK_LAUNCH();
  for (int k = 1; k < g->degree; ++k) {
    int rc = apply_op(g, L, L.d, L.w, st);
    if (rc) return rc;
    const double rho_new = 1.0 / (2.0 * sigma - rho);
    mg_step_kernel<<<L.vgrid, BLK, 0, st>>>(L.n, L.w, L.dinv, rho_new * rho, 2.0 * rho_new / delta, L.r, L.d, x);
    PYN_CHECK_LAUNCH();
    rho = rho_new;
  }
  if (want_resid) {
    int rc = apply_op(g, L, L.d, L.w, st);
    if (rc) return rc;
    mg_resid_kernel<<<L.vgrid, BLK, 0, st>>>(L.n, L.w, L.r, nullptr, nullptr);
    PYN_CHECK_LAUNCH();
  }
  return PYN_OK;
}

int cycle(pyn_mg* g, int l, const double* b, double* x, cudaStream_t st) {
  MgLevel& L = g->L[l];
  if (l == g->nlev - 1) {
    if (g->Ainv) {
      const int64_t thr = L.n * 32;
      mg_dense_kernel<<<(int)((thr + BLK - 1) / BLK), BLK, 0, st>>>((int)L.n, g->Ainv, b, x);
      PYN_CHECK_LAUNCH();
      return PYN_OK;
    }
    return smooth(g, L, b, x, true, false, st);   // no inverse given: the smoother alone (one-level method)
  }
  int rc = smooth(g, L, b, x, true, true, st);
  if (rc) return rc;
  MgLevel& C = g->L[l + 1];
  if (L.halo && C.halo) {   // strip -> strip: R gathers the residual of every local fine node, ghosts included
    rc = pyn::halo_exchange(L.halo, L.r, L.halo_comp, st);
    if (rc) return rc;
  }
  rc = pyn_spmv_planned(L.Rplan, L.Rind, L.Rval, L.r, C.r, PYN_SPMV_SET, 0.0, 0.0, nullptr, st);
  if (rc) return rc;
  if (L.halo && !C.halo) {   // strip -> replicated: every rank restricted its own strip: sum, then all run the same cycle
    PYN_REQUIRE(C.n < 2147483647LL, "coarse level too large for one all-reduce");
    rc = pyn::comm_allreduce_sum(L.comm, C.r, (int)C.n, st);
    if (rc) return rc;
  }
  rc = cycle(g, l + 1, C.r, C.x, st);
  if (rc) return rc;
  if (C.halo) {              // P reads the coarse correction at every local coarse node, ghosts included
    rc = pyn::halo_exchange(C.halo, C.x, C.halo_comp, st);
    if (rc) return rc;
  }
  rc = pyn_spmv_planned(L.Pplan, L.Pind, L.Pval, C.x, L.d, PYN_SPMV_SET, 0.0, 0.0, nullptr, st);
  if (rc) return rc;
  rc = apply_op(g, L, L.d, L.w, st);
  if (rc) return rc;
  mg_resid_kernel<<<L.vgrid, BLK, 0, st>>>(L.n, L.w, L.r, L.d, x);
  PYN_CHECK_LAUNCH();
  return smooth(g, L, b, x, false, false, st);
}

int ensure_work(pyn_mg* g) {
  if (g->pool) return PYN_OK;
  size_t tot = 0;
  for (int l = 0; l < g->nlev; ++l) {
    PYN_REQUIRE(g->L[l].n > 0 && g->L[l].dinv && (g->L[l].tmat || g->L[l].plan), "multigrid level not set");
    PYN_REQUIRE(l == g->nlev - 1 || (g->L[l].Pplan && g->L[l].Rplan), "multigrid transfer not set");
    tot += 4 * (((size_t)g->L[l].n + 1) & ~(size_t)1);
  }
  PYN_CHECK_CUDA(cudaMalloc(&g->pool, sizeof(double) * tot));
  PYN_CHECK_CUDA(cudaMemset(g->pool, 0, sizeof(double) * tot));
  double* p = g->pool;
  for (int l = 0; l < g->nlev; ++l) {
    const size_t ld = ((size_t)g->L[l].n + 1) & ~(size_t)1;
    g->L[l].x = p; g->L[l].r = p + ld; g->L[l].d = p + 2 * ld; g->L[l].w = p + 3 * ld;
    p += 4 * ld;
    g->L[l].vgrid = pyn::grid_for(g->L[l].n, BLK, 8);
  }
  return PYN_OK;
}
}  // namespace

namespace pyn {
int mg_apply(pyn_mg* g, const double* d_r, double* d_z, cudaStream_t st) {
  int rc = ensure_work(g);
  if (rc) return rc;
  return cycle(g, 0, d_r, d_z, st);
}
}  // namespace pyn

extern "C" {

int pyn_mg_create(int nlevels, int degree, double ratio, pyn_mg** out) {
  PYN_REQUIRE(out && nlevels >= 1 && nlevels <= PYN_MG_MAX_LEVELS, "1..16 multigrid levels");
  PYN_REQUIRE(degree >= 1 && degree <= 16 && ratio > 1.0, "Chebyshev degree 1..16, interval ratio > 1");
  pyn_mg* g = new pyn_mg();
  memset(g, 0, sizeof(*g));
  g->nlev = nlevels; g->degree = degree; g->ratio = ratio;
  *out = g;
  return PYN_OK;
}

int pyn_mg_set_level(pyn_mg* g, int l, int64_t n, const pyn_tmat* tmat, const pyn_spmv_plan* plan, const int32_t* d_ind,
                     const double* d_val, const double* d_dinv, double lmax) {
  PYN_REQUIRE(g && l >= 0 && l < g->nlev && n > 0 && d_dinv && lmax > 0.0, "bad arguments");
  PYN_REQUIRE(tmat || (plan && d_ind && d_val), "a level needs an element operator or a planned CSR");
  PYN_REQUIRE(!g->pool, "levels are fixed once the cycle has been applied");
  if (tmat) PYN_REQUIRE(tmat->rdof == tmat->cdof && tmat->topo->nnodes * tmat->rdof == n, "element operator does not match the level");
  if (plan) PYN_REQUIRE(plan->nrows == n, "CSR plan does not match the level");
  MgLevel& L = g->L[l];
  L.n = n; L.tmat = tmat; L.plan = plan; L.ind = d_ind; L.val = d_val; L.dinv = d_dinv; L.lmax = lmax;
  return PYN_OK;
}

int pyn_mg_set_transfer(pyn_mg* g, int l, const pyn_spmv_plan* Pplan, const int32_t* d_Pind, const double* d_Pval,
                        const pyn_spmv_plan* Rplan, const int32_t* d_Rind, const double* d_Rval) {
  PYN_REQUIRE(g && l >= 0 && l + 1 < g->nlev && Pplan && d_Pind && d_Pval && Rplan && d_Rind && d_Rval, "bad arguments");
  MgLevel& L = g->L[l];
  PYN_REQUIRE(L.n > 0 && g->L[l + 1].n > 0, "set both levels before their transfer");
  PYN_REQUIRE(Pplan->nrows == L.n && Rplan->nrows == g->L[l + 1].n, "transfer shapes do not match the levels");
  L.Pplan = Pplan; L.Pind = d_Pind; L.Pval = d_Pval; L.Rplan = Rplan; L.Rind = d_Rind; L.Rval = d_Rval;
  return PYN_OK;
}

int pyn_mg_set_coarse_inverse(pyn_mg* g, const double* d_Ainv) {
  PYN_REQUIRE(g, "mg is NULL");
  PYN_REQUIRE(!d_Ainv || g->L[g->nlev - 1].n < (1 << 15), "coarsest level too large for a dense inverse");
  g->Ainv = d_Ainv;
  return PYN_OK;
}

int pyn_mg_set_level_dist(pyn_mg* g, int l, const pyn_comm* comm, const pyn_halo* halo, int ncomp) {
  PYN_REQUIRE(g && comm && l >= 0 && l < g->nlev, "bad arguments");
  PYN_REQUIRE(comm->world == 1 || (halo && ncomp >= 1 && ncomp <= PYN_HALO_MAX_COMP), "halo plan required for world > 1");
  PYN_REQUIRE(l == 0 || g->L[l - 1].halo || comm->world == 1, "partitioned levels must be the first ones of the hierarchy");
  g->L[l].comm = comm; g->L[l].halo = comm->world > 1 ? halo : nullptr; g->L[l].halo_comp = ncomp;
  return PYN_OK;
}

int pyn_mg_set_dist(pyn_mg* g, const pyn_comm* comm, const pyn_halo* halo, int ncomp) {
  return pyn_mg_set_level_dist(g, 0, comm, halo, ncomp);
}

int pyn_mg_apply(pyn_mg* g, const double* d_r, double* d_z, void* stream) {
  PYN_REQUIRE(g && d_r && d_z && d_r != d_z, "bad arguments");
  return pyn::mg_apply(g, d_r, d_z, pyn::as_stream(stream));
}

int pyn_mg_destroy(pyn_mg* g) {
  if (!g) return PYN_OK;
  cudaFree(g->pool);
  delete g;
  return PYN_OK;
}
}
