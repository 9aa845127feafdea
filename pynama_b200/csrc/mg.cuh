// Geometric p/h-multigrid V-cycle used as the preconditioner of CG (`-pc_type mg`): see mg.cu.
#pragma once
#include "dist.cuh"
#include "spmv.cuh"
#include "tmat.cuh"

constexpr int PYN_MG_MAX_LEVELS = 16;

struct MgLevel {
  int64_t n;                  // dofs of this level
  // operator A_l: element operator when one is attached, else the assembled CSR through its stream plan (all borrowed)
  const pyn_tmat* tmat;
  const pyn_spmv_plan* plan;
  const int* ind;
  const double* val;
  const double* dinv;         // 1 / diag(A_l) on the unknowns of this solve, 0 on prescribed dofs (borrowed)
  double lmax;                // upper bound of the spectrum of D^-1 A_l on the unknowns
  // transfers to level l+1 (absent on the coarsest level; borrowed CSR with their stream plans)
  const pyn_spmv_plan* Pplan; // [n_l x n_{l+1}]  interpolation
  const int* Pind;
  const double* Pval;
  const pyn_spmv_plan* Rplan; // [n_{l+1} x n_l]  its transpose
  const int* Rind;
  const double* Rval;
  double *x, *r, *d, *w;      // owned work vectors (x of level 0 is the caller's z)
  int vgrid;
  // Partitioned level (multi-GPU): the level lives on this rank's strip -- its operator needs the ghost entries of its
  // input refreshed (halo); NULL: the level is replicated (every rank holds and computes all of it)
  const pyn_comm* comm;
  const pyn_halo* halo;
  int halo_comp;
};

struct pyn_mg {
  int nlev, degree;
  double ratio;               // Chebyshev interval [lmax / ratio, lmax]
  MgLevel L[PYN_MG_MAX_LEVELS];
  // Multi-GPU: the first levels (the fine mesh at every order) keep the strip partition; from the first h-coarsened
  // level on the hierarchy is REPLICATED: the restricted residual is summed over the ranks, every rank runs the same
  // coarse cycle on the whole mesh and interpolates to its own strip.
  const double* Ainv;         // dense inverse of the coarsest operator on its unknowns, [n_c][n_c] (borrowed)
  double* pool;               // owned: all work vectors
  int64_t launches_per_apply;
};

namespace pyn {
// z = V(r) on the unknowns, 0 on prescribed dofs; r is not modified.  Pure launches on `st` (graph-capturable).
int mg_apply(pyn_mg* g, const double* d_r, double* d_z, cudaStream_t st);
}  // namespace pyn
