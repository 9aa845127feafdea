// Element handle shared by elem.cu (integration kernels) and assemble.cu (fused integrate-and-assemble).
#pragma once
#include "common.cuh"
#include "elem_affine.cuh"

struct pyn_elem {
  int n, nqF, nqG;
  double alpha_w, alpha_d;
  double* d_tab;  // hF dF wF xiF hG dG wG xiG D wO xiO (device)
  int* d_perm;    // nn
  int off_hF, off_dF, off_wF, off_xF, off_hG, off_dG, off_wG, off_xG, off_D, off_wO, off_xO, tab_len;
  double* d_aff;  // [6][n*n] SF MF CF SG MG CG: 1-D matrices of the affine closed form (elem_affine.cuh)
  double* d_pairK;  // [nn*nn][PYN_AFF_PK] products of the 1-D matrices per local node pair
  double* d_pairR;  // [nn*nn][PYN_AFF_PR]
  unsigned long long* d_count;  // scratch counter (non-affine elements)
};

namespace pyn {
static inline AffTables aff_tables(const pyn_elem* e) {
  const int n2 = e->n * e->n;
  AffTables T;
  T.SF = e->d_aff; T.MF = e->d_aff + n2; T.CF = e->d_aff + 2 * n2;
  T.SG = e->d_aff + 3 * n2; T.MG = e->d_aff + 4 * n2; T.CG = e->d_aff + 5 * n2;
  T.perm = e->d_perm; T.n = e->n; T.alpha_w = e->alpha_w; T.alpha_d = e->alpha_d;
  T.pairK = e->d_pairK; T.pairR = e->d_pairR;
  return T;
}
}  // namespace pyn
