// Krylov solvers on device with PETSc semantics (KSPCG / KSPGMRES + KSPConvergedDefault, left preconditioning,
// preconditioned residual norm, zero or given initial guess) -- replaces KspSolver (src/solver/ksp_solver.py:9-19).
//
// Control flow lives on the device: every kernel of an iteration starts by reading a `done` flag in a small state
// block, so a whole chunk of iterations can be enqueued (as one CUDA graph) without the host knowing whether
// convergence happened in the middle; the host polls the state once per chunk.  The iteration count, the stopping
// iteration and the iterate are therefore exactly those of the sequential algorithm.
// All reductions are two-level with a fixed tree (per-CTA partial -> every CTA re-sums the partials in the same
// order), i.e. bitwise reproducible run to run.
#include <vector>

#include "dist.cuh"
#include "spmv.cuh"
#include "ksp.cuh"

namespace {
constexpr int BLK = 256;



__global__ void jacobi_inv_kernel(int64_t n, const double* __restrict__ diag, double* __restrict__ dinv, int use) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const double d = diag[i];
    dinv[i] = (use && d != 0.0) ? 1.0 / d : 1.0;  // PCJacobi replaces zero diagonal entries by one
  }
}

// r = b, x = 0, z = M^-1 r, partials of z.z and z.r
__global__ void cg_init_kernel(int64_t n, const double* __restrict__ b, CgBuffers B,
                               const uint8_t* __restrict__ owned) {
  __shared__ double sm[BLK / 32];
  double szz = 0.0, szr = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const double r = (owned && !owned[i]) ? 0.0 : b[i];  // rows of ghost nodes belong to another rank
    const double z = B.dinv[i] * r;
    B.r[i] = r;
    B.z[i] = z;
    B.x[i] = 0.0;
    B.p[i] = 0.0;
    szz += z * z;
    szr += z * r;
  }
  szz = pyn::block_sum<BLK>(szz, sm);
  szr = pyn::block_sum<BLK>(szr, sm);
  if (threadIdx.x == 0) {
    B.part_a[blockIdx.x] = szz;
    B.part_b[blockIdx.x] = szr;
  }
}

// partial sums -> scalars (multi-GPU: the scalars are then all-reduced over the ranks)
__global__ void cg_reduce_kernel(int nparts, const double* __restrict__ pa, const double* __restrict__ pb,
                                 double* __restrict__ out) {
  const double a = pyn::reduce_partials(pa, nparts);
  const double b = pb ? pyn::reduce_partials(pb, nparts) : 0.0;
  if (threadIdx.x == 0) {
    out[0] = a;
    if (pb) out[1] = b;
  }
}

// nonzero initial guess (-ksp_initial_guess_nonzero): clear the flags of the previous solve so the product kernels run
__global__ void cg_reset_kernel(KspState* s) {
  s->done = 0;
  s->its = 0;
  s->reason = 0;
  s->pend = 0;
}

// ... then, with p = x0 and w = A p:  r = b - w, x = x0, z = M^-1 r; partials of z.z, z.r and |M^-1 b|^2 (KSPCG:
// KSP_MatMult + VecAYPX; KSPConvergedDefault measures rtol against the preconditioned right-hand side in this case)
__global__ void cg_init_guess_kernel(int64_t n, const double* __restrict__ b, CgBuffers B,
                                     const uint8_t* __restrict__ owned, double* __restrict__ part_c) {
  __shared__ double sm[BLK / 32];
  double szz = 0.0, szr = 0.0, sbb = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const bool mine = !(owned && !owned[i]);
    const double bi = mine ? b[i] : 0.0;
    const double r = mine ? bi - B.w[i] : 0.0;
    const double d = B.dinv[i];
    const double z = d * r, zb = d * bi;
    B.r[i] = r;
    B.z[i] = z;
    B.x[i] = B.p[i];
    szz += z * z;
    szr += z * r;
    sbb += zb * zb;
  }
  szz = pyn::block_sum<BLK>(szz, sm);
  szr = pyn::block_sum<BLK>(szr, sm);
  sbb = pyn::block_sum<BLK>(sbb, sm);
  if (threadIdx.x == 0) {
    B.part_a[blockIdx.x] = szz;
    B.part_b[blockIdx.x] = szr;
    part_c[blockIdx.x] = sbb;
  }
}

// g (multi-GPU only): global sums [p.w, z.z, z.r, |M^-1 b|^2] produced by cg_reduce_kernel + all-reduce
// part_c (nonzero initial guess only): partials of |M^-1 b|^2
__global__ void cg_scalar_init_kernel(int nparts, CgBuffers B, double rtol, double atol, const double* __restrict__ g,
                                      const double* __restrict__ part_c) {
  const double zz = g ? g[1] : pyn::reduce_partials(B.part_a, nparts);
  const double zr = g ? g[2] : pyn::reduce_partials(B.part_b, nparts);
  double bb = -1.0;
  if (part_c) bb = g ? g[3] : pyn::reduce_partials(part_c, nparts);
  if (threadIdx.x == 0) {
    KspState* s = B.st;
    s->dp = sqrt(zz);
    s->rnorm0 = bb > 0.0 ? sqrt(bb) : s->dp;   // zero right-hand side with a nonzero guess: the initial residual
    s->ttol = fmax(rtol * s->rnorm0, atol);
    s->beta = zr;
    s->betaold = 1.0;
    s->its = 0;
    s->reason = 0;
    s->done = 0;
    s->pend = 0;
    s->alpha = 0.0;
    if (s->dp <= s->ttol) {  // KSPConvergedDefault at iteration 0
      s->reason = (s->dp <= atol) ? 3 : 2;
      s->done = 1;
    }
  }
}

// x += alpha_prev p (the previous iteration's step, deferred to here because p is read anyway: one vector pass less
// per iteration than updating x next to r), then p = z + (beta/betaold) p   (first iteration: p = z).
// 16-byte accesses (the work vectors are 16-byte aligned)
__global__ void cg_pupdate_kernel(int64_t n, CgBuffers B) {
  const KspState* s = B.st;
  if (s->done) return;
  const bool first = (s->its == 0);
  const double b = first ? 0.0 : s->beta / s->betaold;
  const double a = s->alpha;
  const int64_t n2 = n >> 1;
  const double2* z2 = reinterpret_cast<const double2*>(B.z);
  double2* p2 = reinterpret_cast<double2*>(B.p);
  double2* x2 = reinterpret_cast<double2*>(B.x);
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n2; i += (int64_t)gridDim.x * BLK) {
    const double2 z = z2[i];
    double2 p = z;
    if (!first) {
      const double2 q = p2[i];
      double2 x = x2[i];
      x.x += a * q.x;
      x.y += a * q.y;
      x2[i] = x;
      p.x = z.x + b * q.x;
      p.y = z.y + b * q.y;
    }
    p2[i] = p;
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const int64_t i = n - 1;
    if (first) {
      B.p[i] = B.z[i];
    } else {
      const double q = B.p[i];
      B.x[i] += a * q;
      B.p[i] = B.z[i] + b * q;
    }
  }
}

// w = A p, per-CTA partial of p.w  (CSR-stream kernel, see spmv.cuh)
struct CgDotEpi {
  double* w;
  const double* p;
  double* dot;
  __device__ __forceinline__ void operator()(int64_t row, double s) const {
    w[row] = s;
    *dot += p[row] * s;
  }
};
__global__ void __launch_bounds__(pyn::SP_BLK, 6)
cg_spmv_dot_kernel(pyn::SpmvPlanView P, const int* __restrict__ ind, const double* __restrict__ val, CgBuffers B) {
  __shared__ pyn::SpmvSmem sm;
  __shared__ double red[pyn::SP_BLK / 32];
  if (B.st->done) return;
  double dot = 0.0;
  CgDotEpi epi{B.w, B.p, &dot};
  for (int b = blockIdx.x; b < P.nblk; b += gridDim.x) pyn::spmv_stream_chunk(P, ind, val, B.p, b, sm, epi);
  dot = pyn::block_sum<pyn::SP_BLK>(dot, red);
  if (threadIdx.x == 0) B.part_a[blockIdx.x] = dot;
}

// bulk-copy (TMA) variant of the same kernel: spmv.cuh spmv_bulk_loop; partial gridDim.x - ... see enqueue_cg_product:
// the last chunk of the matrix is summed by cg_spmv_dot_tail_kernel into the partial after the bulk grid's
__global__ void __launch_bounds__(pyn::SB_THREADS, 4)
cg_spmv_dot_bulk_kernel(pyn::SpmvPlanView P, const int* __restrict__ ind, const double* __restrict__ val, CgBuffers B) {
  extern __shared__ __align__(16) unsigned char sb_raw[];
  __shared__ double red[pyn::SB_BLK / 32];
  if (B.st->done) return;
  pyn::SbSmem& sm = *reinterpret_cast<pyn::SbSmem*>(sb_raw);
  double dot = 0.0;
  CgDotEpi epi{B.w, B.p, &dot};
  pyn::spmv_bulk_loop(P, ind, val, B.p, blockIdx.x, gridDim.x, P.nbblk - 1, sm, epi);
  if (threadIdx.x >= pyn::SB_BLK) return;             // producer warp
  // consumer warps only: named barrier over the 256 consumer threads
  dot = pyn::warp_sum(dot);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) red[w] = dot;
  asm volatile("bar.sync 1, %0;" ::"n"(pyn::SB_BLK) : "memory");
  if (w == 0) {
    double r = (lane < pyn::SB_BLK / 32) ? red[lane] : 0.0;
    r = pyn::warp_sum(r);
    if (lane == 0) B.part_a[blockIdx.x] = r;
  }
}
__global__ void __launch_bounds__(pyn::SP_BLK)
cg_spmv_dot_tail_kernel(pyn::SpmvPlanView P, const int* __restrict__ ind, const double* __restrict__ val, CgBuffers B,
                        int slot) {
  __shared__ pyn::SpmvSmem sm;
  __shared__ double red[pyn::SP_BLK / 32];
  if (B.st->done) return;
  double dot = 0.0;
  CgDotEpi epi{B.w, B.p, &dot};
  pyn::spmv_stream_rows(P, ind, val, B.p, P.brow[P.nbblk - 1], P.brow[P.nbblk], sm, epi);
  dot = pyn::block_sum<pyn::SP_BLK>(dot, red);
  if (threadIdx.x == 0) B.part_a[slot] = dot;
}

// the same product through the element operator (tmat.cuh): phase 1 is pyn::tmat_apply on p; here phase 2 (gather
// per un-flagged node, one thread each) and the flagged rows from the CSR, both with the p.w partials fused
template <int RD>
__global__ void __launch_bounds__(256)
cg_tmat_interior_kernel(pyn::TmatView m, CgBuffers B) {
  __shared__ double red[256 / 32];
  if (B.st->done) return;
  double dot = 0.0;
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < m.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    double acc[RD];
    if (m.flags[r]) {   // row written by cg_tmat_boundary_kernel (launched before): only its share of p.w is due
#pragma unroll
      for (int a = 0; a < RD; ++a) acc[a] = B.w[r * RD + a];
    } else {
      pyn::eop_gather_node<RD>(m, r, acc);
#pragma unroll
      for (int a = 0; a < RD; ++a) B.w[r * RD + a] = acc[a];
    }
#pragma unroll
    for (int a = 0; a < RD; ++a) dot += B.p[r * RD + a] * acc[a];
  }
  dot = pyn::block_sum<256>(dot, red);
  if (threadIdx.x == 0) B.part_a[blockIdx.x] = dot;
}
// flagged rows from the CSR, one warp per row, one row per warp (the list is short: full-width grid, no partials)
__global__ void __launch_bounds__(256)
cg_tmat_boundary_kernel(pyn::TmatView m, int rd, CgBuffers B) {
  if (B.st->done) return;
  const int lane = threadIdx.x & 31;
  const int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (i >= m.nlist * rd) return;
  const int64_t row = (int64_t)m.list[i / rd] * rd + (i % rd);
  const double s = pyn::spmv_row<32>(m.ptr, m.ind, m.val, B.p, row, true, lane);
  if (lane == 0) B.w[row] = s;
}

// a = beta / (p.w) ; r -= a w ; z = M^-1 r ; partials of z.z, z.r   (input partials: nparts_in of p.w);
// x += a p is left to the next cg_pupdate_kernel / cg_finish_kernel (state: alpha, pend)
__global__ void cg_update_kernel(int64_t n, int nparts_in, CgBuffers B, double* __restrict__ out_a,
                                 double* __restrict__ out_b, const double* __restrict__ g) {
  __shared__ double sm[BLK / 32];
  __shared__ double s_a;
  const KspState* s = B.st;
  if (s->done) return;
  if (threadIdx.x < 32) {
    const double dpi = g ? g[0] : pyn::reduce_partials(B.part_a, nparts_in);
    // indefinite operator (p.w <= 0): take no step, so that r, z and x keep the last good iterate (PETSc returns it)
    if (threadIdx.x == 0) s_a = (dpi > 0.0) ? s->beta / dpi : 0.0;
  }
  __syncthreads();
  const double a = s_a;
  double szz = 0.0, szr = 0.0;
  const int64_t n2 = n >> 1;
  double2* r2 = reinterpret_cast<double2*>(B.r);
  double2* z2 = reinterpret_cast<double2*>(B.z);
  const double2* w2 = reinterpret_cast<const double2*>(B.w);
  const double2* d2 = reinterpret_cast<const double2*>(B.dinv);
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n2; i += (int64_t)gridDim.x * BLK) {
    const double2 w = w2[i], d = d2[i];
    double2 r = r2[i], z;
    r.x -= a * w.x;
    r.y -= a * w.y;
    z.x = d.x * r.x;
    z.y = d.y * r.y;
    r2[i] = r;
    z2[i] = z;
    szz += z.x * z.x;
    szr += z.x * r.x;
    szz += z.y * z.y;
    szr += z.y * r.y;
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const int64_t i = n - 1;
    const double r = B.r[i] - a * B.w[i];
    const double z = B.dinv[i] * r;
    B.r[i] = r;
    B.z[i] = z;
    szz += z * z;
    szr += z * r;
  }
  szz = pyn::block_sum<BLK>(szz, sm);
  szr = pyn::block_sum<BLK>(szr, sm);
  if (threadIdx.x == 0) {
    out_a[blockIdx.x] = szz;
    out_b[blockIdx.x] = szr;
  }
}

// ---- CG with a general (multigrid) preconditioner: z = M^-1 r is a sequence of launches between these kernels ----
// r = b (0 on rows owned by another rank), x = 0, p = 0
__global__ void cg_mg_init_kernel(int64_t n, const double* __restrict__ b, CgBuffers B,
                                  const uint8_t* __restrict__ owned) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    B.r[i] = (owned && !owned[i]) ? 0.0 : b[i];
    B.x[i] = 0.0;
    B.p[i] = 0.0;
  }
}
// nonzero initial guess, with p = x0 and w = A p:  r = b - w, x = x0
__global__ void cg_mg_init_guess_kernel(int64_t n, const double* __restrict__ b, CgBuffers B,
                                        const uint8_t* __restrict__ owned) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    B.r[i] = (owned && !owned[i]) ? 0.0 : b[i] - B.w[i];
    B.x[i] = B.p[i];
  }
}
// a = beta / (p.w) ; r -= a w
__global__ void cg_mg_rupdate_kernel(int64_t n, int nparts_in, CgBuffers B, const double* __restrict__ g) {
  __shared__ double s_a;
  const KspState* s = B.st;
  if (s->done) return;
  if (threadIdx.x < 32) {
    const double dpi = g ? g[0] : pyn::reduce_partials(B.part_a, nparts_in);
    if (threadIdx.x == 0) s_a = (dpi > 0.0) ? s->beta / dpi : 0.0;
  }
  __syncthreads();
  const double a = s_a;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) B.r[i] -= a * B.w[i];
}
// the cycle left z on the unknowns (mask != 0) and 0 on prescribed dofs: those rows are identity rows, Jacobi there;
// partials of z.z and z.src  (src = r, or b when |M^-1 b| is wanted)
__global__ void cg_mg_zdot_kernel(int64_t n, const double* __restrict__ mask, const double* __restrict__ src, CgBuffers B,
                                  double* __restrict__ out_a, double* __restrict__ out_b) {
  __shared__ double sm[BLK / 32];
  double szz = 0.0, szr = 0.0;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    const double r = src[i];
    double z = B.z[i];
    if (mask[i] == 0.0) {
      z = B.dinv[i] * r;
      B.z[i] = z;
    }
    szz += z * z;
    szr += z * r;
  }
  szz = pyn::block_sum<BLK>(szz, sm);
  szr = pyn::block_sum<BLK>(szr, sm);
  if (threadIdx.x == 0) {
    out_a[blockIdx.x] = szz;
    if (out_b) out_b[blockIdx.x] = szr;
  }
}

// end of iteration: norms, convergence test, new beta  (one warp)
__global__ void cg_scalar_kernel(int nparts_pw, int nparts, CgBuffers B, const double* __restrict__ pzz,
                                 const double* __restrict__ pzr, double atol, double dtol, int max_it,
                                 const double* __restrict__ g) {
  KspState* s = B.st;
  if (s->done) return;
  const double dpi = g ? g[0] : pyn::reduce_partials(B.part_a, nparts_pw);
  const double zz = g ? g[1] : pyn::reduce_partials(pzz, nparts);
  const double zr = g ? g[2] : pyn::reduce_partials(pzr, nparts);
  if (threadIdx.x == 0) {
    s->its += 1;
    s->alpha = (dpi > 0.0) ? s->beta / dpi : 0.0;   // the same quotient cg_update_kernel used
    s->pend = (dpi > 0.0) ? 1 : 0;
    s->betaold = s->beta;
    s->dp = sqrt(zz);
    if (!(dpi > 0.0)) {  // KSP_DIVERGED_INDEFINITE_MAT
      s->reason = -8;
      s->done = 1;
    } else if (s->dp <= s->ttol) {
      s->reason = (s->dp <= atol) ? 3 : 2;
      s->done = 1;
    } else if (s->dp >= dtol * s->rnorm0) {
      s->reason = -4;
      s->done = 1;
    } else if (s->its >= max_it) {
      s->reason = -3;
      s->done = 1;
    }
    s->beta = zr;
    if (!s->done && zr == 0.0) {  // PETSc: beta == 0 at the top of the next iteration -> CONVERGED_ATOL
      s->its += 1;
      s->reason = 3;
      s->done = 1;
    }
  }
}

// solution out: x plus the step of the last iteration that no p-update has applied yet
__global__ void cg_finish_kernel(int64_t n, CgBuffers B, double* __restrict__ dst) {
  const KspState* s = B.st;
  const bool pend = s->pend != 0;
  const double a = s->alpha;
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) {
    double x = B.x[i];
    if (pend) x += a * B.p[i];
    dst[i] = x;
  }
}

__global__ void copy_masked_kernel(int64_t n, const double* __restrict__ src, double* __restrict__ dst,
                                   const uint8_t* __restrict__ owned) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK)
    dst[i] = (owned && !owned[i]) ? 0.0 : src[i];
}

__global__ void copy_kernel(int64_t n, const double* __restrict__ src, double* __restrict__ dst) {
  for (int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x; i < n; i += (int64_t)gridDim.x * BLK) dst[i] = src[i];
}

}  // namespace

#include "ksp.cuh"


namespace {

// w = A p (+ per-CTA partials of p.w in part_a, *npw of them); ghost entries of p fetched first on N ranks
int enqueue_cg_product(const pyn_ksp* k, cudaStream_t st, int* npw_out) {
  const bool dist = k->comm && k->comm->world > 1;
  if (dist) {  // ghost entries of p <- their owners (MPIAIJ MatMult's VecScatter in the reference)
    int rc = pyn::halo_exchange(k->halo, k->B.p, k->halo_comp, st);
    if (rc) return rc;
  }
  int npw = k->sgrid;   // number of per-CTA partials of p.w
  if (k->tmat) {
    const pyn::TmatView tv = pyn::view_of(k->tmat);
    const int gb = k->tmat->nlist > 0 ? k->tmat->grid_bnd : 0, gi = k->tmat->grid_int;
    int rc = pyn::tmat_apply(k->tmat, k->B.p, &k->B.st->done, st);
    if (rc) return rc;
    if (gb) cg_tmat_boundary_kernel<<<gb, 256, 0, st>>>(tv, k->tmat->rdof, k->B);
    if (k->tmat->rdof == 2) cg_tmat_interior_kernel<2><<<gi, 256, 0, st>>>(tv, k->B);
    else if (k->tmat->rdof == 1) cg_tmat_interior_kernel<1><<<gi, 256, 0, st>>>(tv, k->B);
    else cg_tmat_interior_kernel<3><<<gi, 256, 0, st>>>(tv, k->B);
    npw = gi;
  } else if (k->bulk) {
    const int g = k->plan->bulk_grid;
    cg_spmv_dot_bulk_kernel<<<g, pyn::SB_THREADS, sizeof(pyn::SbSmem), st>>>(pyn::view_of(k->plan), k->ind, k->val, k->B);
    cg_spmv_dot_tail_kernel<<<1, pyn::SP_BLK, 0, st>>>(pyn::view_of(k->plan), k->ind, k->val, k->B, g);
    npw = g + 1;
  } else {
    cg_spmv_dot_kernel<<<k->sgrid, pyn::SP_BLK, 0, st>>>(pyn::view_of(k->plan), k->ind, k->val, k->B);
  }
  *npw_out = npw;
  return PYN_OK;
}

int enqueue_cg_iteration(const pyn_ksp* k, cudaStream_t st) {
  const bool dist = k->comm && k->comm->world > 1;
  const double* g = dist ? k->gsum : nullptr;
  cg_pupdate_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, k->B);
  int npw = 0;
  {
    int rc = enqueue_cg_product(k, st, &npw);
    if (rc) return rc;
  }
  if (dist) {
    int rc = pyn::comm_reduce_allreduce(k->comm, npw, k->B.part_a, 0, nullptr, k->gsum, st);
    if (rc) return rc;
  }
  if (k->mg) {   // z = V-cycle(r): r -= a w, the cycle's launches, then the Jacobi rows + the two dot products
    cg_mg_rupdate_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, npw, k->B, g);
    int rc = pyn::mg_apply(k->mg, k->B.r, k->B.z, st);
    if (rc) return rc;
    cg_mg_zdot_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, k->mg->L[0].dinv, k->B.r, k->B, k->pzz, k->pzr);
  } else
  cg_update_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, npw, k->B, k->pzz, k->pzr, g);
  if (dist) {
    int rc = pyn::comm_reduce_allreduce(k->comm, k->vgrid, k->pzz, k->vgrid, k->pzr, k->gsum + 1, st);
    if (rc) return rc;
  }
  cg_scalar_kernel<<<1, 32, 0, st>>>(npw, k->vgrid, k->B, k->pzz, k->pzr, k->opts.atol, k->opts.dtol,
                                     k->opts.max_it, g);
  return PYN_OK;
}

}  // namespace

extern "C" {

int pyn_ksp_create(int64_t nrows, const int64_t* d_ptr, const int32_t* d_ind, const double* d_val, int64_t nnz,
                   const pyn_ksp_opts* opts, void* stream, pyn_ksp** out) {
  PYN_REQUIRE(nrows > 0 && d_ptr && d_ind && d_val && opts && out, "bad arguments");
  PYN_REQUIRE(opts->type == 0 || opts->type == 1, "ksp type must be 0 (cg) or 1 (gmres)");
  PYN_REQUIRE(opts->pc >= 0 && opts->pc <= 2 && !(opts->pc == 2 && opts->type != 0), "pc must be 0 (none), 1 (jacobi) or 2 (mg, cg only)");
  cudaStream_t st = pyn::as_stream(stream);
  pyn_ksp* k = new pyn_ksp();
  memset(k, 0, sizeof(*k));
  k->n = nrows; k->nnz = nnz; k->ptr = d_ptr; k->ind = d_ind; k->val = d_val; k->opts = *opts;
  if (k->opts.check_every <= 0) k->opts.check_every = opts->pc == 2 ? 1 : 16;   // a V-cycle iteration is ~10^4 x a host poll
  if (k->opts.max_it <= 0) k->opts.max_it = 10000;
  k->vgrid = pyn::grid_for(nrows, BLK, 8);
  {
    int rc0 = pyn_spmv_plan_create(nrows, d_ptr, nnz, stream, &k->plan);
    if (rc0) { delete k; return rc0; }
    int occ = 0;
    PYN_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, cg_spmv_dot_kernel, pyn::SP_BLK, 0));
    const int64_t cap = (int64_t)pyn::sm_count() * (occ > 0 ? occ : 1);
    k->sgrid = (int)(k->plan->nblk < cap ? k->plan->nblk : cap);
    k->bulk = k->plan->bulk_ok && ((reinterpret_cast<uintptr_t>(d_ind) | reinterpret_cast<uintptr_t>(d_val)) & 15) == 0;
    if (k->bulk)
      PYN_CHECK_CUDA(cudaFuncSetAttribute(cg_spmv_dot_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)sizeof(pyn::SbSmem)));
  }
  double* pool = nullptr;
  size_t nparts = (size_t)(k->vgrid > k->sgrid ? k->vgrid : k->sgrid);
  if (nparts < (size_t)k->plan->bulk_grid + 1) nparts = (size_t)k->plan->bulk_grid + 1;
  if (nparts < (size_t)pyn::sm_count() * 16) nparts = (size_t)pyn::sm_count() * 16;   // room for the template operator's grids
  const size_t ld = ((size_t)nrows + 1) & ~(size_t)1;   // even stride: every work vector stays 16-byte aligned
  PYN_CHECK_CUDA(cudaMalloc(&pool, sizeof(double) * (6 * ld + 4 * nparts)));
  k->B.r = pool; k->B.z = pool + ld; k->B.p = pool + 2 * ld; k->B.w = pool + 3 * ld;
  k->B.x = pool + 4 * ld; k->B.dinv = pool + 5 * ld;
  k->B.part_a = pool + 6 * ld; k->B.part_b = k->B.part_a + nparts;
  k->pzz = k->B.part_b + nparts; k->pzr = k->pzz + nparts;
  PYN_CHECK_CUDA(cudaMalloc(&k->B.st, sizeof(KspState)));
  PYN_CHECK_CUDA(cudaMallocHost(&k->h_state, sizeof(KspState)));
  // Jacobi: dinv = 1/diag(A)
  int rc = pyn_csr_diagonal(nrows, d_ptr, d_ind, d_val, k->B.w, stream);
  if (rc) return rc;
  jacobi_inv_kernel<<<k->vgrid, BLK, 0, st>>>(nrows, k->B.w, k->B.dinv, opts->pc >= 1);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  *out = k;
  return PYN_OK;
}

int pyn_ksp_solve(pyn_ksp* k, const double* d_b, double* d_x, pyn_ksp_result* res, void* stream) {
  PYN_REQUIRE(k && d_b && d_x, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  if (k->opts.type == 1) return pyn::gmres_solve(k, d_b, d_x, res, st);
  PYN_REQUIRE(k->opts.pc != 2 || k->mg, "pc mg: attach the hierarchy first (pyn_ksp_set_mg)");
  int64_t launches = 0;
  const bool dist = k->comm && k->comm->world > 1;
  const bool guess = k->opts.nonzero_guess != 0;
  if (k->mg && dist) PYN_REQUIRE(k->mg->L[0].comm == k->comm, "multigrid hierarchy of a partitioned mesh: attach the communicator (pyn_mg_set_dist)");
  if (guess) {   // r0 = b - A x0 with the contents of d_x (KSPSetInitialGuessNonzero)
    cg_reset_kernel<<<1, 1, 0, st>>>(k->B.st);
    copy_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, d_x, k->B.p);
    int npw = 0;
    int rc = enqueue_cg_product(k, st, &npw);
    if (rc) return rc;
    if (k->mg) {   // |M^-1 b| first (KSPConvergedDefault's reference norm), then z = M^-1 r0
      copy_masked_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, d_b, k->B.r, dist ? k->owned : nullptr);   // b on owned rows
      rc = pyn::mg_apply(k->mg, k->B.r, k->B.z, st);
      if (rc) return rc;
      cg_mg_zdot_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, k->mg->L[0].dinv, k->B.r, k->B, k->pzz, nullptr);
      cg_mg_init_guess_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, d_b, k->B, dist ? k->owned : nullptr);
      rc = pyn::mg_apply(k->mg, k->B.r, k->B.z, st);
      if (rc) return rc;
      cg_mg_zdot_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, k->mg->L[0].dinv, k->B.r, k->B, k->B.part_a, k->B.part_b);
    } else
    cg_init_guess_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, d_b, k->B, dist ? k->owned : nullptr, k->pzz);
    launches += 4;
  } else if (k->mg) {
    cg_mg_init_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, d_b, k->B, dist ? k->owned : nullptr);
    int rc = pyn::mg_apply(k->mg, k->B.r, k->B.z, st);
    if (rc) return rc;
    cg_mg_zdot_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, k->mg->L[0].dinv, k->B.r, k->B, k->B.part_a, k->B.part_b);
  } else {
    cg_init_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, d_b, k->B, dist ? k->owned : nullptr);
  }
  PYN_CHECK_LAUNCH();
  if (dist) {
    cg_reduce_kernel<<<1, 32, 0, st>>>(k->vgrid, k->B.part_a, k->B.part_b, k->gsum + 1);
    if (guess) cg_reduce_kernel<<<1, 32, 0, st>>>(k->vgrid, k->pzz, nullptr, k->gsum + 3);
    PYN_CHECK_LAUNCH();
    int rc = pyn::comm_allreduce_sum(k->comm, k->gsum + 1, guess ? 3 : 2, st);
    if (rc) return rc;
  }
  cg_scalar_init_kernel<<<1, 32, 0, st>>>(k->vgrid, k->B, k->opts.rtol, k->opts.atol, dist ? k->gsum : nullptr,
                                          guess ? k->pzz : nullptr);
  PYN_CHECK_LAUNCH();
  launches += 2;
  const int chunk = k->opts.check_every;
  if (!k->graph || k->graph_iters != chunk) {
    if (k->graph) cudaGraphExecDestroy(k->graph);
    k->graph = nullptr;
    cudaStream_t cs;
    PYN_CHECK_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    cudaGraph_t g;
    PYN_CHECK_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
    int erc = PYN_OK;
    for (int i = 0; i < chunk && erc == PYN_OK; ++i) erc = enqueue_cg_iteration(k, cs);
    if (erc != PYN_OK) {   // keep the message of the launch that failed, not the one of the aborted capture
      cudaStreamEndCapture(cs, &g);
      cudaGetLastError();
      cudaStreamDestroy(cs);
      return erc;
    }
    PYN_CHECK_CUDA(cudaStreamEndCapture(cs, &g));
    PYN_CHECK_CUDA(cudaGraphInstantiate(&k->graph, g, 0));
    PYN_CHECK_CUDA(cudaGraphDestroy(g));
    PYN_CHECK_CUDA(cudaStreamDestroy(cs));
    k->graph_iters = chunk;
  }
  int enq = 0;
  for (;;) {
    PYN_CHECK_CUDA(cudaMemcpyAsync(k->h_state, k->B.st, sizeof(KspState), cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaStreamSynchronize(st));
    if (k->h_state->done || enq >= k->opts.max_it) break;
    PYN_CHECK_CUDA(cudaGraphLaunch(k->graph, st));
    enq += chunk;
    launches += 4 * (int64_t)chunk;
  }
  pyn::count_launch((int)((dist ? 6 : 4) * (int64_t)k->h_state->its));  // only the launches that did work are claimed
  cg_finish_kernel<<<k->vgrid, BLK, 0, st>>>(k->n, k->B, d_x);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  if (res) {
    res->its = k->h_state->its;
    res->reason = k->h_state->reason ? k->h_state->reason : -3;
    res->rnorm0 = k->h_state->rnorm0;
    res->rnorm = k->h_state->dp;
  }
  k->last_spmv = k->h_state->its;
  k->last_launches = launches + 1;
  return PYN_OK;
}

int pyn_ksp_destroy(pyn_ksp* k) {
  if (!k) return PYN_OK;
  if (k->graph) cudaGraphExecDestroy(k->graph);
  pyn_spmv_plan_destroy(k->plan);
  cudaFree(k->B.r);
  cudaFree(k->B.st);
  cudaFree(k->gsum);
  pyn::gmres_free(k->gm);
  cudaFreeHost(k->h_state);
  delete k;
  return PYN_OK;
}

int pyn_ksp_set_dist(pyn_ksp* k, const pyn_comm* comm, const pyn_halo* halo, int ncomp, const uint8_t* d_owned_rows) {
  PYN_REQUIRE(k && comm, "bad arguments");
  PYN_REQUIRE(comm->world == 1 || (halo && ncomp >= 1 && ncomp <= PYN_HALO_MAX_COMP), "halo plan required for world > 1");
  k->comm = comm; k->halo = halo; k->halo_comp = ncomp; k->owned = d_owned_rows;
  if (!k->gsum) PYN_CHECK_CUDA(cudaMalloc(&k->gsum, 4 * sizeof(double)));
  if (k->graph) { cudaGraphExecDestroy(k->graph); k->graph = nullptr; }  // re-capture with the collectives
  return PYN_OK;
}

int pyn_ksp_set_tmat(pyn_ksp* k, const pyn_tmat* m) {
  PYN_REQUIRE(k, "ksp is NULL");
  if (m) {
    PYN_REQUIRE(m->rdof == m->cdof && m->topo->nnodes * m->rdof == k->n, "template operator does not match the system");
  }
  k->tmat = m;
  if (k->graph) { cudaGraphExecDestroy(k->graph); k->graph = nullptr; }
  return PYN_OK;
}

int pyn_ksp_set_mg(pyn_ksp* k, pyn_mg* g) {
  PYN_REQUIRE(k, "ksp is NULL");
  PYN_REQUIRE(k->opts.type == 0, "the multigrid preconditioner is available for cg");
  if (g) PYN_REQUIRE(g->L[0].n == k->n, "multigrid hierarchy does not match the system");
  k->mg = g;
  if (k->graph) { cudaGraphExecDestroy(k->graph); k->graph = nullptr; }
  return PYN_OK;
}

int pyn_ksp_stats(const pyn_ksp* k, int64_t* spmv, int64_t* launches) {
  PYN_REQUIRE(k, "ksp is NULL");
  if (spmv) *spmv = k->last_spmv;
  if (launches) *launches = k->last_launches;
  return PYN_OK;
}
}
