// Closed-form element matrices on AFFINE quadrilaterals (parallelograms: every box / stretched / sheared mesh of the
// reference), shared by the integration kernels (elem.cu) and the fused integrate-and-assemble kernels (assemble.cu).
//
// On an affine element the Jacobian of the bilinear 4-corner map is constant, so the per-point geometric factors of
// Spectral.getElemKLEMatrices (src/elements/spectral.py:89-157) leave the quadrature sums and every block of K, Rw, Rd
// becomes a short combination of Kronecker products of six 1-D matrices (SURVEY 8a', here for a general parallelogram):
//     M^R[a,c] = sum_q w_q h_a(q) h_c(q)    S^R[a,c] = sum_q w_q h_a'(q) h_c'(q)    C^R[a,c] = sum_q w_q h_a(q) h_c'(q)
// for the rule R in {F = full, G = reduced}, with 17 per-element coefficients (13 for K, 4 for Rw / Rd).  It is the SAME
// quadrature as the general kernel, summed in a different order (<= 1e-14 relative apart); what makes it worth having is
// that an entry costs ~6 multiply-adds instead of a contraction over the quadrature points, so the element matrix
// never has to exist in memory: the assembly kernels evaluate the entries they need while they write the CSR rows.
//
// Every operation is an explicit round-to-nearest intrinsic: the materialising kernel (elem_kle_affine_kernel) and the
// fused assembly kernels therefore produce bit-identical entries whatever the compiler does around the inlined code.
#pragma once
#include "common.cuh"

constexpr int PYN_AFF_NCOEF = 18;   // per element: 13 (K) + 4 (Rw, Rd) + 1 pad (16-byte rows)

namespace pyn {

struct AffTables {   // device pointers (or shared-memory copies) of the six n x n matrices + the local -> tensor map
  const double *SF, *MF, *CF, *SG, *MG, *CG;
  const int* perm;   // [nn]: element-local node -> a*n+b
  int n;
  double alpha_w, alpha_d;
  const double* pairK;   // [nn*nn][PYN_AFF_PK] product records (device)
  const double* pairR;   // [nn*nn][PYN_AFF_PR]
};

// coefficients of one element from its corner coordinates (closure order); returns false when the element is not
// affine (relative bilinear term above 1e-13: such elements take the general sum-factorised kernel)
__device__ __forceinline__ bool affine_coefficients(const double* __restrict__ c, double aw, double ad,
                                                    double* __restrict__ out) {
  // local corner k sits at (r,s) = (+1,+1), (-1,+1), (-1,-1), (+1,-1)   (spectral.py:226)
  const double xr = 0.25 * (c[0] - c[2] - c[4] + c[6]), xs = 0.25 * (c[0] + c[2] - c[4] - c[6]);
  const double yr = 0.25 * (c[1] - c[3] - c[5] + c[7]), ys = 0.25 * (c[1] + c[3] - c[5] - c[7]);
  const double bx = 0.25 * (c[0] - c[2] + c[4] - c[6]), by = 0.25 * (c[1] - c[3] + c[5] - c[7]);
  const double size = fabs(xr) + fabs(xs) + fabs(yr) + fabs(ys);
  const double det = xr * ys - yr * xs;
  const double inv = 1.0 / det;
  const double mxr = ys * inv, mxs = -yr * inv, myr = -xs * inv, mys = xr * inv;
  out[0] = det * (mxr * mxr + myr * myr);            // grad.grad  rr | rs | ss
  out[1] = det * (mxr * mxs + myr * mys);
  out[2] = det * (mxs * mxs + mys * mys);
  out[3] = det * (ad * mxr * mxr + aw * myr * myr);  // uu: ad x x^T + aw y y^T
  out[4] = det * (ad * mxr * mxs + aw * myr * mys);
  out[5] = det * (ad * mxs * mxs + aw * mys * mys);
  out[6] = det * (ad * myr * myr + aw * mxr * mxr);  // vv: ad y y^T + aw x x^T
  out[7] = det * (ad * myr * mys + aw * mxr * mxs);
  out[8] = det * (ad * mys * mys + aw * mxs * mxs);
  out[9] = det * (ad * mxr * myr - aw * myr * mxr);  // uv: ad x y^T - aw y x^T   rr | rs | sr | ss
  out[10] = det * (ad * mxr * mys - aw * myr * mxs);
  out[11] = det * (ad * mxs * myr - aw * mys * mxr);
  out[12] = det * (ad * mxs * mys - aw * mys * mxs);
  out[13] = det * mxr;                               // first-order blocks (Rw, Rd)
  out[14] = det * mxs;
  out[15] = det * myr;
  out[16] = det * mys;
  out[17] = 0.0;
  return (fabs(bx) + fabs(by)) <= 1e-13 * size;
}

__device__ __forceinline__ double aff_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double aff_fma(double a, double b, double c) { return __fma_rn(a, b, c); }

constexpr int PYN_AFF_PK = 8;   // doubles per local node pair in the K product table (7 used; 64-byte records)
constexpr int PYN_AFF_PR = 4;   // ... in the Rw / Rd product table

// The element-independent half of an entry: products of the 1-D matrices for the local node pair (i, j).  They are
// tabulated once per element order (pyn_elem.d_pairK / d_pairR, [nn*nn] records) so that an assembly thread evaluates an
// entry from one 64-byte (32-byte) record and the element's coefficients -- no index arithmetic, no divisions.
__device__ __forceinline__ void affine_pair_products_K(const AffTables& T, int i, int j, double* __restrict__ pr) {
  const int n = T.n, pi = T.perm[i], pj = T.perm[j];
  const int a = pi / n, b = pi - a * n, c = pj / n, d = pj - c * n;
  const int ac = a * n + c, ca = c * n + a, bd = b * n + d, db = d * n + b;
  pr[0] = aff_mul(T.SF[ac], T.MF[bd]);
  pr[1] = __dadd_rn(aff_mul(T.CF[ca], T.CF[bd]), aff_mul(T.CF[ac], T.CF[db]));   // un-fused pair: bitwise symmetric
  pr[2] = aff_mul(T.MF[ac], T.SF[bd]);
  pr[3] = aff_mul(T.SG[ac], T.MG[bd]);
  pr[4] = aff_mul(T.CG[ca], T.CG[bd]);
  pr[5] = aff_mul(T.CG[ac], T.CG[db]);
  pr[6] = aff_mul(T.MG[ac], T.SG[bd]);
  pr[7] = 0.0;
}
__device__ __forceinline__ void affine_pair_products_R(const AffTables& T, int i, int j, double* __restrict__ pr) {
  const int n = T.n, pi = T.perm[i], pj = T.perm[j];
  const int a = pi / n, b = pi - a * n, c = pj / n, d = pj - c * n;
  const int ac = a * n + c, ca = c * n + a, bd = b * n + d, db = d * n + b;
  pr[0] = aff_mul(T.CF[ac], T.MF[bd]);   // test phi, trial gradient
  pr[1] = aff_mul(T.MF[ac], T.CF[bd]);
  pr[2] = aff_mul(T.CG[ca], T.MG[bd]);   // test gradient, trial phi
  pr[3] = aff_mul(T.MG[ac], T.CG[db]);
}

// 2 x 2 block of K_e between two local nodes from their product record: [[uu, uv], [vu, vv]]
__device__ __forceinline__ void affine_block_K(const double* __restrict__ cf, const double* __restrict__ pr,
                                               double (&v)[2][2]) {
  // full rule: the vector Laplacian part, identical on uu and vv
  double gr = aff_mul(cf[0], pr[0]);
  gr = aff_fma(cf[1], pr[1], gr);
  gr = aff_fma(cf[2], pr[2], gr);
  // reduced rule: penalty terms
  const double sm = pr[3], c1 = pr[4], c2 = pr[5], ms = pr[6];
  const double cs = __dadd_rn(c1, c2);
  double kuu = aff_mul(cf[3], sm);
  kuu = aff_fma(cf[4], cs, kuu);
  kuu = aff_fma(cf[5], ms, kuu);
  double kvv = aff_mul(cf[6], sm);
  kvv = aff_fma(cf[7], cs, kvv);
  kvv = aff_fma(cf[8], ms, kvv);
  // uv and vu = uv^T: the same four products with the two mixed coefficients exchanged; the mixed pair is summed
  // un-fused so that K_e[i][j] (uv) and K_e[j][i] (vu seen from the other node) carry the same bits
  double kuv = aff_mul(cf[9], sm);
  kuv = __dadd_rn(kuv, __dadd_rn(aff_mul(cf[10], c1), aff_mul(cf[11], c2)));
  kuv = aff_fma(cf[12], ms, kuv);
  double kvu = aff_mul(cf[9], sm);
  kvu = __dadd_rn(kvu, __dadd_rn(aff_mul(cf[11], c1), aff_mul(cf[10], c2)));
  kvu = aff_fma(cf[12], ms, kvu);
  v[0][0] = __dadd_rn(gr, kuu);
  v[0][1] = kuv;
  v[1][0] = kvu;
  v[1][1] = __dadd_rn(gr, kvv);
}

// 2 x 1 block of Rw_e (which == 1) or Rd_e (which == 2); cf = the element's coefficients 13..16
__device__ __forceinline__ void affine_block_R(const double* __restrict__ cf, const double* __restrict__ pr, int which,
                                               double alpha_w, double alpha_d, double (&v)[2][1]) {
  const double Ax = aff_fma(cf[1], pr[1], aff_mul(cf[0], pr[0])), Ay = aff_fma(cf[3], pr[1], aff_mul(cf[2], pr[0]));
  const double Cx = aff_fma(cf[1], pr[3], aff_mul(cf[0], pr[2])), Cy = aff_fma(cf[3], pr[3], aff_mul(cf[2], pr[2]));
  if (which == 1) {        // Rw = int dv . curl(w) + alpha_w int curl(dv) w      (spectral.py:140-147)
    v[0][0] = aff_fma(-alpha_w, Cy, Ay);
    v[1][0] = aff_fma(alpha_w, Cx, -Ax);
  } else {                 // Rd = - int dv . grad(D) + alpha_d int div(dv) D
    v[0][0] = aff_fma(alpha_d, Cx, -Ax);
    v[1][0] = aff_fma(alpha_d, Cy, -Ay);
  }
}

// Row a of the same blocks, entry by entry the same operations as affine_block_K / affine_block_R (same bits): the
// signature-tile assembly evaluates the two rows of a node at DIFFERENT column positions when it aligns its stores.
__device__ __forceinline__ void affine_row_K(const double* __restrict__ cf, const double* __restrict__ pr, int a,
                                             double (&v)[2]) {
  double gr = aff_mul(cf[0], pr[0]);
  gr = aff_fma(cf[1], pr[1], gr);
  gr = aff_fma(cf[2], pr[2], gr);
  const double sm = pr[3], c1 = pr[4], c2 = pr[5], ms = pr[6];
  const double cs = __dadd_rn(c1, c2);
  if (a == 0) {
    double kuu = aff_mul(cf[3], sm);
    kuu = aff_fma(cf[4], cs, kuu);
    kuu = aff_fma(cf[5], ms, kuu);
    double kuv = aff_mul(cf[9], sm);
    kuv = __dadd_rn(kuv, __dadd_rn(aff_mul(cf[10], c1), aff_mul(cf[11], c2)));
    kuv = aff_fma(cf[12], ms, kuv);
    v[0] = __dadd_rn(gr, kuu);
    v[1] = kuv;
  } else {
    double kvv = aff_mul(cf[6], sm);
    kvv = aff_fma(cf[7], cs, kvv);
    kvv = aff_fma(cf[8], ms, kvv);
    double kvu = aff_mul(cf[9], sm);
    kvu = __dadd_rn(kvu, __dadd_rn(aff_mul(cf[11], c1), aff_mul(cf[10], c2)));
    kvu = aff_fma(cf[12], ms, kvu);
    v[0] = kvu;
    v[1] = __dadd_rn(gr, kvv);
  }
}
__device__ __forceinline__ double affine_row_R(const double* __restrict__ cf, const double* __restrict__ pr, int which,
                                               double alpha_w, double alpha_d, int a) {
  // the x pair feeds row 1 of Rw and row 0 of Rd, the y pair the other two
  const bool usex = (which == 1) == (a == 1);
  const double c0 = usex ? cf[0] : cf[2], c1 = usex ? cf[1] : cf[3];
  const double A = aff_fma(c1, pr[1], aff_mul(c0, pr[0])), Cc = aff_fma(c1, pr[3], aff_mul(c0, pr[2]));
  if (which == 1) return a == 0 ? aff_fma(-alpha_w, Cc, A) : aff_fma(alpha_w, Cc, -A);
  return aff_fma(alpha_d, Cc, -A);
}

}  // namespace pyn
