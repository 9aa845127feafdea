// Batched element integration on straight-sided quads (dim = 2), sum-factorised.
//
// What it computes: Spectral.getElemKLEMatrices / getElemKLEOperators of the reference
// (src/elements/spectral.py:89-157, :159-218).  How: instead of one dense (2nn x 2nn) rank update per
// quadrature point (O(n^6) per element) every block of K/Rw/Rd is written as
//     B[(a,b),(c,d)] = sum_q1 X[q1,a] X'[q1,c] ( sum_q2 g(q1,q2) Y[q2,b] Y'[q2,d] )
// with X,Y in {h, h'} the 1-D Lagrange tables and g the per-point geometric factor of the bilinear 4-corner map:
// step 1 builds T[q1][b][d] in shared memory (O(n^4)), step 2 contracts q1 (O(n^5)).
//
// One CTA per element (grid-stride).  All FP64.
#include "common.cuh"
#include "elem.cuh"

namespace {

struct Geo {
  double ax, bx, cx, ay, by, cy;
};

__device__ __forceinline__ Geo make_geo(const double* __restrict__ c) {
  // local corner k sits at (r,s) = (+1,+1), (-1,+1), (-1,-1), (+1,-1)   (spectral.py:226, 2-node tabulation)
  Geo g;
  g.ax = 0.25 * (c[0] - c[2] - c[4] + c[6]);
  g.cx = 0.25 * (c[0] + c[2] - c[4] - c[6]);
  g.bx = 0.25 * (c[0] - c[2] + c[4] - c[6]);
  g.ay = 0.25 * (c[1] - c[3] - c[5] + c[7]);
  g.cy = 0.25 * (c[1] + c[3] - c[5] - c[7]);
  g.by = 0.25 * (c[1] - c[3] + c[5] - c[7]);
  return g;
}

// J = [[x_r, y_r],[x_s, y_s]] ; inv(J) rows: m_x = (y_s, -y_r)/det , m_y = (-x_s, x_r)/det
struct PointGeo {
  double det, mxr, mxs, myr, mys;
};
__device__ __forceinline__ PointGeo point_geo(const Geo& g, double r, double s) {
  const double xr = g.ax + g.bx * s, yr = g.ay + g.by * s;
  const double xs = g.cx + g.bx * r, ys = g.cy + g.by * r;
  PointGeo p;
  p.det = xr * ys - yr * xs;
  const double inv = 1.0 / p.det;
  p.mxr = ys * inv;
  p.mxs = -yr * inv;
  p.myr = -xs * inv;
  p.mys = xr * inv;
  return p;
}

// a*b + c*d without fusion: identical bits when the two products are exchanged
__device__ __forceinline__ double sym_pair(double a, double b, double c, double d) {
  return __dadd_rn(__dmul_rn(a, b), __dmul_rn(c, d));
}

// T[q1][b][d] = sum_q2 c[q1][q2] * Ya[q2][b] * Yb[q2][d]
__device__ __forceinline__ void build_T(double* __restrict__ T, const double* __restrict__ c,
                                        const double* __restrict__ Ya, const double* __restrict__ Yb, int nq, int n) {
  const int tot = nq * n * n;
  for (int idx = threadIdx.x; idx < tot; idx += blockDim.x) {
    const int d = idx % n, b = (idx / n) % n, q1 = idx / (n * n);
    double acc = 0.0;
    for (int q2 = 0; q2 < nq; ++q2) acc += c[q1 * nq + q2] * (Ya[q2 * n + b] * Yb[q2 * n + d]);
    T[idx] = acc;
  }
}

__global__ void __launch_bounds__(256)
elem_kle_kernel(pyn_elem E, const double* __restrict__ coords, int64_t ne, double* __restrict__ Kout,
                double* __restrict__ Rwout, double* __restrict__ Rdout) {
  extern __shared__ double sm[];
  const int n = E.n, nn = n * n, nqF = E.nqF, nqG = E.nqG;
  double* tab = sm;                       // tab_len doubles
  int* perm = (int*)(tab + E.tab_len);    // nn ints (padded to even)
  double* fld = tab + E.tab_len + (nn + 1) / 2;
  const int nF2 = nqF * nqF, nG2 = nqG * nqG;
  double* fF = fld;                       // 7 fields on the full rule
  double* fG = fF + 7 * nF2;              // 14 fields on the reduced rule
  double* Tbuf = fG + 14 * nG2;
  const int tF = nqF * nn, tG = nqG * nn; // one T array

  for (int i = threadIdx.x; i < E.tab_len; i += blockDim.x) tab[i] = E.d_tab[i];
  for (int i = threadIdx.x; i < nn; i += blockDim.x) perm[i] = E.d_perm[i];
  __syncthreads();
  const double *hF = tab + E.off_hF, *dF = tab + E.off_dF, *wF = tab + E.off_wF, *xF = tab + E.off_xF;
  const double *hG = tab + E.off_hG, *dG = tab + E.off_dG, *wG = tab + E.off_wG, *xG = tab + E.off_xG;
  const double aw = E.alpha_w, ad = E.alpha_d;

  for (int64_t e = blockIdx.x; e < ne; e += gridDim.x) {
    const Geo g = make_geo(coords + e * 8);
    // ---- per-point geometric fields -----------------------------------------------------------------
    for (int idx = threadIdx.x; idx < nF2; idx += blockDim.x) {
      const int q1 = idx / nqF, q2 = idx % nqF;
      const PointGeo p = point_geo(g, xF[q1], xF[q2]);
      const double wd = wF[q1] * wF[q2] * p.det;
      fF[0 * nF2 + idx] = wd * (p.mxr * p.mxr + p.myr * p.myr);  // grad.grad  rr
      fF[1 * nF2 + idx] = wd * (p.mxr * p.mxs + p.myr * p.mys);  //            rs
      fF[2 * nF2 + idx] = wd * (p.mxs * p.mxs + p.mys * p.mys);  //            ss
      fF[3 * nF2 + idx] = wd * p.mxr;
      fF[4 * nF2 + idx] = wd * p.mxs;
      fF[5 * nF2 + idx] = wd * p.myr;
      fF[6 * nF2 + idx] = wd * p.mys;
    }
    for (int idx = threadIdx.x; idx < nG2; idx += blockDim.x) {
      const int q1 = idx / nqG, q2 = idx % nqG;
      const PointGeo p = point_geo(g, xG[q1], xG[q2]);
      const double wd = wG[q1] * wG[q2] * p.det;
      const double xx_rr = p.mxr * p.mxr, xx_rs = p.mxr * p.mxs, xx_ss = p.mxs * p.mxs;
      const double yy_rr = p.myr * p.myr, yy_rs = p.myr * p.mys, yy_ss = p.mys * p.mys;
      // uu: ad*x x^T + aw*y y^T ; vv: ad*y y^T + aw*x x^T ; uv: ad*x y^T - aw*y x^T
      fG[0 * nG2 + idx] = wd * (ad * xx_rr + aw * yy_rr);
      fG[1 * nG2 + idx] = wd * (ad * xx_rs + aw * yy_rs);
      fG[2 * nG2 + idx] = wd * (ad * xx_ss + aw * yy_ss);
      fG[3 * nG2 + idx] = wd * (ad * yy_rr + aw * xx_rr);
      fG[4 * nG2 + idx] = wd * (ad * yy_rs + aw * xx_rs);
      fG[5 * nG2 + idx] = wd * (ad * yy_ss + aw * xx_ss);
      fG[6 * nG2 + idx] = wd * (ad * p.mxr * p.myr - aw * p.myr * p.mxr);  // rr
      fG[7 * nG2 + idx] = wd * (ad * p.mxr * p.mys - aw * p.myr * p.mxs);  // rs
      fG[8 * nG2 + idx] = wd * (ad * p.mxs * p.myr - aw * p.mys * p.mxr);  // sr
      fG[9 * nG2 + idx] = wd * (ad * p.mxs * p.mys - aw * p.mys * p.mxs);  // ss
      fG[10 * nG2 + idx] = wd * p.mxr;
      fG[11 * nG2 + idx] = wd * p.mxs;
      fG[12 * nG2 + idx] = wd * p.myr;
      fG[13 * nG2 + idx] = wd * p.mys;
    }
    __syncthreads();

    if (Kout != nullptr) {
      // ---- phase K, step 1: 4 T arrays on the full rule, 12 on the reduced rule ------------------------
      double* TF = Tbuf;             // 4 x tF
      double* TG = Tbuf + 4 * tF;    // 12 x tG
      // phi_r: X=d,Y=h ; phi_s: X=h,Y=d
      build_T(TF + 0 * tF, fF + 0 * nF2, hF, hF, nqF, n);
      build_T(TF + 1 * tF, fF + 1 * nF2, hF, dF, nqF, n);
      build_T(TF + 2 * tF, fF + 1 * nF2, dF, hF, nqF, n);
      build_T(TF + 3 * tF, fF + 2 * nF2, dF, dF, nqF, n);
#pragma unroll 1
      for (int blk = 0; blk < 2; ++blk) {  // uu, vv (symmetric coefficient)
        build_T(TG + (4 * blk + 0) * tG, fG + (3 * blk + 0) * nG2, hG, hG, nqG, n);
        build_T(TG + (4 * blk + 1) * tG, fG + (3 * blk + 1) * nG2, hG, dG, nqG, n);
        build_T(TG + (4 * blk + 2) * tG, fG + (3 * blk + 1) * nG2, dG, hG, nqG, n);
        build_T(TG + (4 * blk + 3) * tG, fG + (3 * blk + 2) * nG2, dG, dG, nqG, n);
      }
      build_T(TG + 8 * tG, fG + 6 * nG2, hG, hG, nqG, n);   // uv rr
      build_T(TG + 9 * tG, fG + 7 * nG2, hG, dG, nqG, n);   // uv rs
      build_T(TG + 10 * tG, fG + 8 * nG2, dG, hG, nqG, n);  // uv sr
      build_T(TG + 11 * tG, fG + 9 * nG2, dG, dG, nqG, n);  // uv ss
      __syncthreads();
      // ---- phase K, step 2 -----------------------------------------------------------------------------
      double* Ke = Kout + e * (int64_t)(4 * nn) * nn;
      for (int idx = threadIdx.x; idx < nn * nn; idx += blockDim.x) {
        const int i = idx / nn, j = idx % nn;
        const int pi = perm[i], pj = perm[j];
        const int a = pi / n, b = pi % n, c = pj / n, d = pj % n;
        const int bd = b * n + d, db = d * n + b;
        double gr = 0.0, kuu = 0.0, kvv = 0.0, kuv = 0.0, kvu = 0.0;
        for (int q = 0; q < nqF; ++q) {
          const double da = dF[q * n + a], ha = hF[q * n + a], dc = dF[q * n + c], hc = hF[q * n + c];
          const double* t = TF + q * nn + bd;
          // the two mixed terms are added as an un-fused pair so that K comes out bitwise symmetric
          gr = fma(da * dc, t[0], gr);
          gr += sym_pair(da * hc, t[tF], ha * dc, t[2 * tF]);
          gr = fma(ha * hc, t[3 * tF], gr);
        }
        for (int q = 0; q < nqG; ++q) {
          const double da = dG[q * n + a], ha = hG[q * n + a], dc = dG[q * n + c], hc = hG[q * n + c];
          const double x11 = da * dc, x12 = da * hc, x21 = ha * dc, x22 = ha * hc;
          const double* t = TG + q * nn + bd;
          const double* tt = TG + q * nn + db;
          kuu = fma(x11, t[0], kuu);
          kuu += sym_pair(x12, t[tG], x21, t[2 * tG]);
          kuu = fma(x22, t[3 * tG], kuu);
          kvv = fma(x11, t[4 * tG], kvv);
          kvv += sym_pair(x12, t[5 * tG], x21, t[6 * tG]);
          kvv = fma(x22, t[7 * tG], kvv);
          kuv = fma(x11, t[8 * tG], kuv);
          kuv += sym_pair(x12, t[9 * tG], x21, t[10 * tG]);
          kuv = fma(x22, t[11 * tG], kuv);
          kvu = fma(x11, tt[8 * tG], kvu);
          kvu += sym_pair(x12, tt[10 * tG], x21, tt[9 * tG]);
          kvu = fma(x22, tt[11 * tG], kvu);
        }
        double* row0 = Ke + (int64_t)(2 * i) * (2 * nn) + 2 * j;
        double* row1 = row0 + 2 * nn;
        *reinterpret_cast<double2*>(row0) = make_double2(gr + kuu, kuv);
        *reinterpret_cast<double2*>(row1) = make_double2(kvu, gr + kvv);
      }
      __syncthreads();
    }

    if (Rwout != nullptr || Rdout != nullptr) {
      // ---- phase R: A_x,A_y on the full rule (test phi, trial grad), C_x,C_y on the reduced rule -------
      double* TF = Tbuf;           // 4 x tF : (mx_r: h,h) (mx_s: h,d) (my_r: h,h) (my_s: h,d)
      double* TG = Tbuf + 4 * tF;  // 4 x tG : (mx_r: h,h) (mx_s: d,h) (my_r: h,h) (my_s: d,h)
      build_T(TF + 0 * tF, fF + 3 * nF2, hF, hF, nqF, n);
      build_T(TF + 1 * tF, fF + 4 * nF2, hF, dF, nqF, n);
      build_T(TF + 2 * tF, fF + 5 * nF2, hF, hF, nqF, n);
      build_T(TF + 3 * tF, fF + 6 * nF2, hF, dF, nqF, n);
      build_T(TG + 0 * tG, fG + 10 * nG2, hG, hG, nqG, n);
      build_T(TG + 1 * tG, fG + 11 * nG2, dG, hG, nqG, n);
      build_T(TG + 2 * tG, fG + 12 * nG2, hG, hG, nqG, n);
      build_T(TG + 3 * tG, fG + 13 * nG2, dG, hG, nqG, n);
      __syncthreads();
      double* Rwe = Rwout ? Rwout + e * (int64_t)(2 * nn) * nn : nullptr;
      double* Rde = Rdout ? Rdout + e * (int64_t)(2 * nn) * nn : nullptr;
      for (int idx = threadIdx.x; idx < nn * nn; idx += blockDim.x) {
        const int i = idx / nn, j = idx % nn;
        const int pi = perm[i], pj = perm[j];
        const int a = pi / n, b = pi % n, c = pj / n, d = pj % n;
        const int bd = b * n + d;
        double Ax = 0.0, Ay = 0.0, Cx = 0.0, Cy = 0.0;
        for (int q = 0; q < nqF; ++q) {
          const double ha = hF[q * n + a], dc = dF[q * n + c], hc = hF[q * n + c];
          const double* t = TF + q * nn + bd;
          Ax += (ha * dc) * t[0] + (ha * hc) * t[tF];
          Ay += (ha * dc) * t[2 * tF] + (ha * hc) * t[3 * tF];
        }
        for (int q = 0; q < nqG; ++q) {
          const double da = dG[q * n + a], ha = hG[q * n + a], hc = hG[q * n + c];
          const double* t = TG + q * nn + bd;
          Cx += (da * hc) * t[0] + (ha * hc) * t[tG];
          Cy += (da * hc) * t[2 * tG] + (ha * hc) * t[3 * tG];
        }
        if (Rwe) {
          Rwe[(int64_t)(2 * i) * nn + j] = Ay - aw * Cy;
          Rwe[(int64_t)(2 * i + 1) * nn + j] = -Ax + aw * Cx;
        }
        if (Rde) {
          Rde[(int64_t)(2 * i) * nn + j] = -Ax + ad * Cx;
          Rde[(int64_t)(2 * i + 1) * nn + j] = -Ay + ad * Cy;
        }
      }
      __syncthreads();
    }
  }
}

// Nodal-quadrature operators: the shape functions are the identity at the nodes, so every (point k, node j) pair
// gives closed-form 3x2 / 2x3 / 1x2 blocks (spectral.py:181-217).
__global__ void __launch_bounds__(256)
elem_ops_kernel(pyn_elem E, const double* __restrict__ coords, int64_t ne, double* __restrict__ SrT,
                double* __restrict__ Div, double* __restrict__ Curl, double* __restrict__ W) {
  extern __shared__ double sm[];
  const int n = E.n, nn = n * n;
  double* tab = sm;
  int* perm = (int*)(tab + E.tab_len);
  for (int i = threadIdx.x; i < E.tab_len; i += blockDim.x) tab[i] = E.d_tab[i];
  for (int i = threadIdx.x; i < nn; i += blockDim.x) perm[i] = E.d_perm[i];
  __syncthreads();
  const double *D = tab + E.off_D, *wO = tab + E.off_wO, *xO = tab + E.off_xO;
  for (int64_t e = blockIdx.x; e < ne; e += gridDim.x) {
    const Geo g = make_geo(coords + e * 8);
    double* S = SrT ? SrT + e * (int64_t)(3 * nn) * (2 * nn) : nullptr;
    double* Dv = Div ? Div + e * (int64_t)(2 * nn) * (3 * nn) : nullptr;
    double* C = Curl ? Curl + e * (int64_t)nn * (2 * nn) : nullptr;
    for (int idx = threadIdx.x; idx < nn * nn; idx += blockDim.x) {
      const int k = idx / nn, j = idx % nn;
      const int pk = perm[k], pj = perm[j];
      const int q1 = pk / n, q2 = pk % n, c = pj / n, d = pj % n;
      const PointGeo p = point_geo(g, xO[q1], xO[q2]);
      const double wd = wO[q1] * wO[q2] * p.det;
      const double fr = (q2 == d) ? D[q1 * n + c] : 0.0;  // d phi_j / dr at point k
      const double fs = (q1 == c) ? D[q2 * n + d] : 0.0;
      const double gx = p.mxr * fr + p.mxs * fs, gy = p.myr * fr + p.mys * fs;
      if (W && j == 0) W[e * nn + k] = wd;
      if (C) {
        double* r = C + (int64_t)k * (2 * nn) + 2 * j;
        r[0] = wd * (-gy);
        r[1] = wd * gx;
      }
      if (S) {
        double* r = S + (int64_t)(3 * k) * (2 * nn) + 2 * j;
        r[0] = wd * (0.5 * gx);
        r[1] = wd * (0.5 * -gy);
        r += 2 * nn;
        r[0] = wd * (0.5 * gy);
        r[1] = wd * (0.5 * gx);
        r += 2 * nn;
        r[0] = wd * (0.5 * -gx);
        r[1] = wd * (0.5 * gy);
      }
      if (Dv) {
        double* r = Dv + (int64_t)(2 * k) * (3 * nn) + 3 * j;
        r[0] = wd * gx;
        r[1] = wd * gy;
        r[2] = 0.0;
        r += 3 * nn;
        r[0] = 0.0;
        r[1] = wd * gx;
        r[2] = wd * gy;
      }
    }
  }
}

// ---- affine elements: closed form (elem_affine.cuh) ---------------------------------------------------------------
// One thread per element; corners in and coefficients out go through shared memory so that both global streams are
// coalesced (64 B and 144 B per element: read and written directly they touch a sector per lane and instruction).
constexpr int AFC_T = 128;                   // elements (= threads) per CTA
constexpr int AFC_SI = 9, AFC_SO = PYN_AFF_NCOEF + 1;   // padded strides: 2-way bank conflicts at most
__global__ void __launch_bounds__(AFC_T)
elem_affine_coef_kernel(int64_t ne, const double* __restrict__ coords, double aw, double ad, double* __restrict__ coef,
                        unsigned long long* __restrict__ nonaffine) {
  __shared__ double s_in[AFC_T * AFC_SI], s_out[AFC_T * AFC_SO];
  const int tid = threadIdx.x;
  const int64_t e0 = blockIdx.x * (int64_t)AFC_T;
  const int nb = (int)min((int64_t)AFC_T, ne - e0);
  for (int i = tid; i < nb * 8; i += AFC_T) s_in[(i >> 3) * AFC_SI + (i & 7)] = coords[e0 * 8 + i];
  __syncthreads();
  if (tid < nb) {
    double c[8], out[PYN_AFF_NCOEF];
#pragma unroll
    for (int k = 0; k < 8; ++k) c[k] = s_in[tid * AFC_SI + k];
    const bool ok = pyn::affine_coefficients(c, aw, ad, out);
#pragma unroll
    for (int k = 0; k < PYN_AFF_NCOEF; ++k) s_out[tid * AFC_SO + k] = out[k];
    if (!ok) atomicAdd(nonaffine, 1ull);
  }
  __syncthreads();
  for (int i = tid; i < nb * PYN_AFF_NCOEF; i += AFC_T) {
    const int e = i / PYN_AFF_NCOEF;
    coef[e0 * PYN_AFF_NCOEF + i] = s_out[e * AFC_SO + (i - e * PYN_AFF_NCOEF)];
  }
}

// product records of every local node pair (once per element order)
__global__ void elem_pair_tables_kernel(pyn::AffTables T, double* __restrict__ pairK, double* __restrict__ pairR) {
  const int nn = T.n * T.n;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < nn * nn; idx += gridDim.x * blockDim.x) {
    const int i = idx / nn, j = idx - i * nn;
    pyn::affine_pair_products_K(T, i, j, pairK + (int64_t)idx * pyn::PYN_AFF_PK);
    pyn::affine_pair_products_R(T, i, j, pairR + (int64_t)idx * pyn::PYN_AFF_PR);
  }
}

// materialised K_e / Rw_e / Rd_e from the coefficients: one thread per (element, local node pair)
__global__ void __launch_bounds__(256)
elem_kle_affine_kernel(pyn::AffTables T, const double* __restrict__ coef, int64_t ne, double* __restrict__ Kout,
                       double* __restrict__ Rwout, double* __restrict__ Rdout) {
  const int nn = T.n * T.n;
  const int64_t tot = ne * (int64_t)nn * nn;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < tot; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t e = idx / (nn * nn);
    const int ij = (int)(idx - e * (int64_t)nn * nn), i = ij / nn, j = ij - i * nn;
    const double* cf = coef + e * PYN_AFF_NCOEF;
    if (Kout) {
      double v[2][2];
      pyn::affine_block_K(cf, T.pairK + (int64_t)ij * pyn::PYN_AFF_PK, v);
      double* row0 = Kout + e * (int64_t)(4 * nn) * nn + (int64_t)(2 * i) * (2 * nn) + 2 * j;
      *reinterpret_cast<double2*>(row0) = make_double2(v[0][0], v[0][1]);
      *reinterpret_cast<double2*>(row0 + 2 * nn) = make_double2(v[1][0], v[1][1]);
    }
    if (Rwout) {
      double v[2][1];
      pyn::affine_block_R(cf + 13, T.pairR + (int64_t)ij * pyn::PYN_AFF_PR, 1, T.alpha_w, T.alpha_d, v);
      double* o = Rwout + e * (int64_t)(2 * nn) * nn + (int64_t)(2 * i) * nn + j;
      o[0] = v[0][0];
      o[nn] = v[1][0];
    }
    if (Rdout) {
      double v[2][1];
      pyn::affine_block_R(cf + 13, T.pairR + (int64_t)ij * pyn::PYN_AFF_PR, 2, T.alpha_w, T.alpha_d, v);
      double* o = Rdout + e * (int64_t)(2 * nn) * nn + (int64_t)(2 * i) * nn + j;
      o[0] = v[0][0];
      o[nn] = v[1][0];
    }
  }
}

}  // namespace

extern "C" {

int pyn_elem_create(int ngl, int nqF, const double* hF, const double* dF, const double* wF, const double* xF,
                    int nqG, const double* hG, const double* dG, const double* wG, const double* xG,
                    const double* Dn, const double* wO, const double* xO, const int32_t* perm, double alpha_w,
                    double alpha_d, pyn_elem** out) {
  PYN_REQUIRE(out != nullptr, "out is NULL");
  if (ngl < 2 || ngl > PYN_MAX_NGL) {
    pyn::set_error("pyn_elem_create: ngl=%d outside [2,%d]", ngl, PYN_MAX_NGL);
    return PYN_ERR_LIMIT;
  }
  PYN_REQUIRE(nqF >= 1 && nqF <= PYN_MAX_NGL && nqG >= 1 && nqG <= PYN_MAX_NGL, "bad quadrature size");
  pyn_elem* e = new pyn_elem();
  memset(e, 0, sizeof(*e));
  e->n = ngl;
  e->nqF = nqF;
  e->nqG = nqG;
  e->alpha_w = alpha_w;
  e->alpha_d = alpha_d;
  const int n = ngl;
  int off = 0;
  auto take = [&](int len) { int o = off; off += len; return o; };
  e->off_hF = take(nqF * n); e->off_dF = take(nqF * n); e->off_wF = take(nqF); e->off_xF = take(nqF);
  e->off_hG = take(nqG * n); e->off_dG = take(nqG * n); e->off_wG = take(nqG); e->off_xG = take(nqG);
  e->off_D = take(n * n); e->off_wO = take(n); e->off_xO = take(n);
  e->tab_len = off;
  double* host = new double[off];
  memcpy(host + e->off_hF, hF, sizeof(double) * nqF * n); memcpy(host + e->off_dF, dF, sizeof(double) * nqF * n);
  memcpy(host + e->off_wF, wF, sizeof(double) * nqF);     memcpy(host + e->off_xF, xF, sizeof(double) * nqF);
  memcpy(host + e->off_hG, hG, sizeof(double) * nqG * n); memcpy(host + e->off_dG, dG, sizeof(double) * nqG * n);
  memcpy(host + e->off_wG, wG, sizeof(double) * nqG);     memcpy(host + e->off_xG, xG, sizeof(double) * nqG);
  memcpy(host + e->off_D, Dn, sizeof(double) * n * n);
  memcpy(host + e->off_wO, wO, sizeof(double) * n);       memcpy(host + e->off_xO, xO, sizeof(double) * n);
  cudaError_t err = cudaMalloc(&e->d_tab, sizeof(double) * off);
  if (err == cudaSuccess) err = cudaMemcpy(e->d_tab, host, sizeof(double) * off, cudaMemcpyHostToDevice);
  delete[] host;
  if (err == cudaSuccess) err = cudaMalloc(&e->d_perm, sizeof(int) * n * n);
  if (err == cudaSuccess) err = cudaMemcpy(e->d_perm, perm, sizeof(int) * n * n, cudaMemcpyHostToDevice);
  {  // the six 1-D matrices of the affine closed form (elem_affine.cuh): S, M symmetric by construction
    const int n2 = n * n;
    double* a = new double[6 * n2];
    auto fill = [&](double* S, double* M, double* C, int nq, const double* h, const double* d, const double* w) {
      for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
          double cs = 0.0;
          for (int q = 0; q < nq; ++q) cs += w[q] * h[q * n + i] * d[q * n + j];
          C[i * n + j] = cs;
          if (j < i) continue;
          double ss = 0.0, ms = 0.0;
          for (int q = 0; q < nq; ++q) {
            ss += w[q] * d[q * n + i] * d[q * n + j];
            ms += w[q] * h[q * n + i] * h[q * n + j];
          }
          S[i * n + j] = S[j * n + i] = ss;
          M[i * n + j] = M[j * n + i] = ms;
        }
    };
    fill(a, a + n2, a + 2 * n2, nqF, hF, dF, wF);
    fill(a + 3 * n2, a + 4 * n2, a + 5 * n2, nqG, hG, dG, wG);
    if (err == cudaSuccess) err = cudaMalloc(&e->d_aff, sizeof(double) * 6 * n2);
    if (err == cudaSuccess) err = cudaMemcpy(e->d_aff, a, sizeof(double) * 6 * n2, cudaMemcpyHostToDevice);
    delete[] a;
  }
  if (err == cudaSuccess) err = cudaMalloc(&e->d_count, sizeof(unsigned long long));
  {
    const size_t np = (size_t)n * n * n * n;
    if (err == cudaSuccess) err = cudaMalloc(&e->d_pairK, sizeof(double) * np * pyn::PYN_AFF_PK);
    if (err == cudaSuccess) err = cudaMalloc(&e->d_pairR, sizeof(double) * np * pyn::PYN_AFF_PR);
    if (err == cudaSuccess) {
      elem_pair_tables_kernel<<<(int)((np + 255) / 256), 256>>>(pyn::aff_tables(e), e->d_pairK, e->d_pairR);
      err = cudaDeviceSynchronize();
    }
  }
  if (err != cudaSuccess) {
    pyn::set_error("pyn_elem_create: %s", cudaGetErrorString(err));
    delete e;
    return PYN_ERR_CUDA;
  }
  *out = e;
  return PYN_OK;
}

int pyn_elem_destroy(pyn_elem* e) {
  if (!e) return PYN_OK;
  cudaFree(e->d_tab);
  cudaFree(e->d_perm);
  cudaFree(e->d_aff);
  cudaFree(e->d_pairK);
  cudaFree(e->d_pairR);
  cudaFree(e->d_count);
  delete e;
  return PYN_OK;
}

static size_t kle_smem_bytes(const pyn_elem* e) {
  const int n = e->n, nn = n * n;
  size_t dbl = e->tab_len + (nn + 1) / 2 + 7 * e->nqF * e->nqF + 14 * e->nqG * e->nqG;
  dbl += 4 * (size_t)e->nqF * nn + 12 * (size_t)e->nqG * nn;  // phase K is the larger one
  return dbl * sizeof(double);
}

int pyn_elem_kle(const pyn_elem* e, const double* d_coords, int64_t ne, double* d_K, double* d_Rw, double* d_Rd,
                 void* stream) {
  PYN_REQUIRE(e && d_coords && ne >= 0, "bad arguments");
  if (ne == 0) return PYN_OK;
  const size_t smem = kle_smem_bytes(e);
  if (smem > 227 * 1024) {
    pyn::set_error("pyn_elem_kle: ngl=%d needs %zu B of shared memory (> 227 KB)", e->n, smem);
    return PYN_ERR_LIMIT;
  }
  PYN_CHECK_CUDA(cudaFuncSetAttribute(elem_kle_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // one CTA per element, sized to the element: nn^2 (i, j) pairs of work per phase -- 32 threads at ngl 2 ... 256 from
  // ngl 4 on -- and as many CTAs per SM as threads and shared memory allow (ngl 2: 32, ngl 5: 8)
  const int nn2 = e->n * e->n * e->n * e->n;
  int threads = ((nn2 + 31) / 32) * 32;
  if (threads > 256) threads = 256;
  int ctas_per_sm = (int)((227 * 1024) / (smem + 1024));
  if (ctas_per_sm > 2048 / threads) ctas_per_sm = 2048 / threads;
  if (ctas_per_sm > 32) ctas_per_sm = 32;
  if (ctas_per_sm < 1) ctas_per_sm = 1;
  int64_t grid = (int64_t)pyn::sm_count() * ctas_per_sm;
  if (grid > ne) grid = ne;
  elem_kle_kernel<<<(int)grid, threads, smem, pyn::as_stream(stream)>>>(*e, d_coords, ne, d_K, d_Rw, d_Rd);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_elem_ops(const pyn_elem* e, const double* d_coords, int64_t ne, double* d_SrT, double* d_Div, double* d_Curl,
                 double* d_W, void* stream) {
  PYN_REQUIRE(e && d_coords && ne >= 0, "bad arguments");
  if (ne == 0) return PYN_OK;
  const int nn = e->n * e->n;
  const size_t smem = (e->tab_len + (nn + 1) / 2) * sizeof(double);
  int grid = pyn::sm_count() * 8;
  if (grid > ne) grid = (int)ne;
  elem_ops_kernel<<<grid, 256, smem, pyn::as_stream(stream)>>>(*e, d_coords, ne, d_SrT, d_Div, d_Curl, d_W);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_elem_affine_coef(const pyn_elem* e, const double* d_coords, int64_t ne, double* d_coef, int64_t* h_nonaffine,
                         void* stream) {
  PYN_REQUIRE(e && d_coords && d_coef && ne >= 0, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  if (h_nonaffine) *h_nonaffine = 0;
  if (ne == 0) return PYN_OK;
  PYN_CHECK_CUDA(cudaMemsetAsync(e->d_count, 0, sizeof(unsigned long long), st));
  elem_affine_coef_kernel<<<(unsigned)((ne + AFC_T - 1) / AFC_T), AFC_T, 0, st>>>(ne, d_coords, e->alpha_w, e->alpha_d,
                                                                                    d_coef, e->d_count);
  PYN_CHECK_LAUNCH();
  if (h_nonaffine) {   // the caller decides the path from this count: one scalar read-back
    unsigned long long c = 0;
    PYN_CHECK_CUDA(cudaMemcpyAsync(&c, e->d_count, sizeof(c), cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaStreamSynchronize(st));
    *h_nonaffine = (int64_t)c;
  }
  return PYN_OK;
}

int pyn_elem_kle_affine(const pyn_elem* e, const double* d_coef, int64_t ne, double* d_K, double* d_Rw, double* d_Rd,
                        void* stream) {
  PYN_REQUIRE(e && d_coef && ne >= 0, "bad arguments");
  if (ne == 0) return PYN_OK;
  const int nn = e->n * e->n;
  const int64_t tot = ne * (int64_t)nn * nn;
  elem_kle_affine_kernel<<<pyn::grid_for(tot, 256, 16), 256, 0, pyn::as_stream(stream)>>>(pyn::aff_tables(e), d_coef, ne, d_K,
                                                                                       d_Rw, d_Rd);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}
}
