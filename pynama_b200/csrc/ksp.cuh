// Krylov solver handle shared by ksp.cu (CG, handle life cycle) and gmres.cu.
#pragma once
#include "dist.cuh"
#include "spmv.cuh"
#include "tmat.cuh"
#include "mg.cuh"

struct KspState {
  double beta, betaold, dp, rnorm0, ttol;
  double alpha;   // CG: step length of the last iteration; x += alpha p is applied by the NEXT p-update (or the finish)
  int its, reason, done, pend;   // pend: that update of x is still outstanding
};

struct CgBuffers {
  double *r, *z, *p, *w, *x, *dinv;
  double *part_a, *part_b;  // per-CTA partials (two reductions per kernel)
  KspState* st;
};

struct pyn_ksp {
  int64_t n, nnz;
  const int64_t* ptr;
  const int* ind;
  const double* val;
  pyn_ksp_opts opts;
  CgBuffers B;
  double *pzz, *pzr;
  KspState* h_state;  // pinned
  int vgrid, sgrid;
  int bulk;   // w = A p through the bulk-copy (TMA) CSR stream
  pyn_spmv_plan* plan;
  cudaGraphExec_t graph;
  int graph_iters;
  int64_t last_spmv, last_launches;
  // multi-GPU (pyn_ksp_set_dist): communicator, halo plan of the unknown (ncomp dofs per node), owned-row mask
  const pyn_comm* comm;
  const pyn_halo* halo;
  int halo_comp;
  const uint8_t* owned;
  double* gsum;  // [p.w, z.z, z.r]
  struct pyn_gmres* gm;  // GMRES work space (lazy)
  const pyn_tmat* tmat;  // template operator used for w = A p instead of the CSR stream (pyn_ksp_set_tmat)
  pyn_mg* mg;            // multigrid preconditioner (pc == 2, pyn_ksp_set_mg); borrowed
};

struct pyn_gmres;
namespace pyn {
int gmres_solve(pyn_ksp* k, const double* d_b, double* d_x, pyn_ksp_result* res, cudaStream_t st);
void gmres_free(pyn_gmres* g);
}  // namespace pyn
