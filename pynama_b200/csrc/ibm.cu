// Immersed-boundary interpolation: the Lagrangian-Eulerian coupling of the velocity correction step.
//
// Reference: ImmersedBoundaryStatic.collectNodes / computeClose / computeDirac / buildIBMMatrix
// (src/cases/immersed_boundary.py:222-251, 258-277, 318-338) and the discrete delta functions of
// src/domain/immersed_body.py:259-264, 398-421.  The reference loops over the Lagrangian nodes in Python and, for the
// dynamic case, does so again every time step (rebuildMatrix, :416-433).  Here both steps are kernels:
//
//   ibm_collect_kernel  one warp per (Lagrangian node, chunk of the candidate Euler nodes, ascending node id) keeps the
//                       close ones with their delta value: ballot + prefix inside the warp, prefix sums of the chunk
//                       counts between them, so the lists come out sorted whatever the launch geometry;
//   ibm_gram_kernel     the Lagrangian system matrix A = H S (S = H^T dl h, H scaled by h^2): one warp per (row node,
//                       32 partner nodes), one partner per lane, a two-pointer merge of the two sorted lists -- terms added in
//                       ascending Euler node order (what a row-times-column product over sorted CSR rows does).
//
// Both run as count / fill pairs on the caller's prefix sums.  Every floating-point operation is an explicit
// round-to-nearest intrinsic in the order the reference's Python expressions evaluate, so the oracle (numpy) and the
// kernels agree to the last bit up to the square root and division, which IEEE rounds correctly on both sides.
#include "common.cuh"

namespace {

// discrete delta functions phi(r), r = |distance| / h >= 0    (immersed_body.py:398-421)
__device__ __forceinline__ double ibm_phi(int kind, double r) {
  if (kind == 0) {   // fourGrid (the reference's default, immersed_body.py:136)
    if (r <= 1.0) {
      const double s = __dsqrt_rn(__dsub_rn(__dadd_rn(1.0, __dmul_rn(4.0, r)), __dmul_rn(4.0, __dmul_rn(r, r))));
      return __ddiv_rn(__dadd_rn(__dsub_rn(3.0, __dmul_rn(2.0, r)), s), 8.0);
    }
    if (r <= 2.0) {
      const double s =
          __dsqrt_rn(__dsub_rn(__dadd_rn(-7.0, __dmul_rn(12.0, r)), __dmul_rn(4.0, __dmul_rn(r, r))));
      return __ddiv_rn(__dsub_rn(__dsub_rn(5.0, __dmul_rn(2.0, r)), s), 8.0);
    }
    return 0.0;
  }
  if (kind == 1) {   // threeGrid
    if (r <= 0.5) {
      const double s = __dsqrt_rn(__dadd_rn(__dmul_rn(-3.0, __dmul_rn(r, r)), 1.0));
      return __ddiv_rn(__dadd_rn(1.0, s), 3.0);
    }
    if (r <= 1.5) {
      const double q = __dsub_rn(1.0, r);
      const double s = __dsqrt_rn(__dadd_rn(__dmul_rn(-3.0, __dmul_rn(q, q)), 1.0));
      return __ddiv_rn(__dsub_rn(__dsub_rn(5.0, __dmul_rn(3.0, r)), s), 6.0);
    }
    return 0.0;
  }
  return r < 1.0 ? __dsub_rn(1.0, r) : 0.0;   // linear
}

// dirac = prod_r phi(|r| / h) / h    (immersed_body.py:259-264: multiply, then divide, per direction)
__device__ __forceinline__ double ibm_dirac(int kind, double dx, double dy, double h) {
  double d = 1.0;
  d = __dmul_rn(d, ibm_phi(kind, __ddiv_rn(fabs(dx), h)));
  d = __ddiv_rn(d, h);
  d = __dmul_rn(d, ibm_phi(kind, __ddiv_rn(fabs(dy), h)));
  d = __ddiv_rn(d, h);
  return d;
}

// mode 0: ImmersedBoundaryStatic.computeClose (:325-336), mode 1: ImmersedBoundaryDynamic.computeClose (:391-402)
__device__ __forceinline__ bool ibm_close(int mode, double dx, double dy, double dl) {
  const double ex = __ddiv_rn(dx, dl), ey = __ddiv_rn(dy, dl);
  if (mode == 0) return fabs(ex) < 3.0 && fabs(ey) < 3.0;
  return fabs(ex) < 2.0 && fabs(__dmul_rn(ey, dl)) < __dadd_rn(1.0, __dmul_rn(2.0, dl));
}

// One warp per (marker, chunk of the candidate list): the counts of the chunks, marker-major, are prefix-summed by the
// caller, which gives every warp the place of its (ascending) piece of the marker's list.
template <bool FILL>
__global__ void __launch_bounds__(256)
ibm_collect_kernel(int64_t nlag, const double* __restrict__ lagxy, int64_t ncand, const int* __restrict__ cand,
                   const double* __restrict__ nodexy, double dl, double h, int mode, int kind, int nchunks,
                   int64_t chunk, const int64_t* __restrict__ ptr, int* __restrict__ count_or_nodes,
                   double* __restrict__ dirac) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (wid >= nlag * nchunks) return;
  const int64_t l = wid / nchunks, ck = wid - l * nchunks;
  const double lx = lagxy[2 * l], ly = lagxy[2 * l + 1];
  const int64_t cbeg = ck * chunk, cend = min(ncand, cbeg + chunk);
  int64_t at = FILL ? ptr[wid] : 0;
  int found = 0;
  for (int64_t c0 = cbeg; c0 < cend; c0 += 32) {
    const int64_t c = c0 + lane;
    int node = 0;
    double dx = 0.0, dy = 0.0;
    bool close = false;
    if (c < cend) {
      node = cand[c];
      dx = __dsub_rn(nodexy[2 * (int64_t)node], lx);        // coords - coordBodyNode
      dy = __dsub_rn(nodexy[2 * (int64_t)node + 1], ly);
      close = ibm_close(mode, dx, dy, dl);
    }
    const unsigned m = __ballot_sync(0xffffffffu, close);
    if (FILL && close) {
      const int64_t o = at + __popc(m & ((1u << lane) - 1u));
      count_or_nodes[o] = node;
      dirac[o] = ibm_dirac(kind, dx, dy, h);
    }
    at += __popc(m);
    found += __popc(m);
  }
  if (!FILL && lane == 0) count_or_nodes[wid] = found;
}

// A = H S restricted to one velocity component (the two components carry the same numbers on the rows 2i+d / columns
// 2j+d): entry (i, j) exists when the node lists of i and j intersect; its value is sum_e (d_i(e) a) (d_j(e) b).
template <bool FILL>
__global__ void __launch_bounds__(256)
ibm_gram_kernel(int64_t nlag, const int64_t* __restrict__ ptr, const int* __restrict__ nodes,
                const double* __restrict__ dirac, double a, double b, int nblk, const int64_t* __restrict__ aptr,
                int* __restrict__ count_or_cols, double* __restrict__ avals) {
  const int lane = threadIdx.x & 31;
  const int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;   // (row marker i, block of 32 partners)
  if (wid >= nlag * nblk) return;
  const int64_t i = wid / nblk, j = (wid - i * nblk) * 32 + lane;
  const int64_t i0 = ptr[i], i1 = ptr[i + 1];
  bool hit = false;
  double s = 0.0;
  if (j < nlag) {
    int64_t p = i0, q = ptr[j];
    const int64_t q1 = ptr[j + 1];
    while (p < i1 && q < q1) {
      const int np_ = nodes[p], nq = nodes[q];
      if (np_ == nq) {
        hit = true;
        s = __dadd_rn(s, __dmul_rn(__dmul_rn(dirac[p], a), __dmul_rn(dirac[q], b)));
        ++p;
        ++q;
      } else if (np_ < nq) {
        ++p;
      } else {
        ++q;
      }
    }
  }
  const unsigned m = __ballot_sync(0xffffffffu, hit);
  if (FILL && hit) {
    const int64_t o = aptr[wid] + __popc(m & ((1u << lane) - 1u));
    count_or_cols[o] = (int)j;
    avals[o] = s;
  }
  if (!FILL && lane == 0) count_or_cols[wid] = __popc(m);
}

}  // namespace

extern "C" {

int pyn_ibm_chunks(int64_t n, int64_t* chunk) {
  // pieces a list of n candidates is cut into (one warp each per marker): about 4096 candidates, whole warps
  int64_t nchunks = (n + 4095) / 4096;
  if (nchunks < 1) nchunks = 1;
  int64_t c = (n + nchunks - 1) / nchunks;
  c = (c + 31) / 32 * 32;
  if (c < 32) c = 32;
  if (chunk) *chunk = c;
  return (int)((n + c - 1) / c > 0 ? (n + c - 1) / c : 1);
}

int pyn_ibm_collect(int64_t nlag, const double* d_lagxy, int64_t ncand, const int32_t* d_cand, const double* d_nodexy,
                    double dl, double h, int mode, int delta_kind, const int64_t* d_ptr, int32_t* d_count_or_nodes,
                    double* d_dirac, void* stream) {
  PYN_REQUIRE(nlag >= 0 && ncand >= 0 && d_lagxy && d_nodexy && d_count_or_nodes, "bad arguments");
  PYN_REQUIRE(dl > 0.0 && h > 0.0 && (mode == 0 || mode == 1) && delta_kind >= 0 && delta_kind <= 2, "bad parameters");
  PYN_REQUIRE(ncand == 0 || d_cand, "candidate list missing");
  PYN_REQUIRE(!d_ptr || d_dirac, "fill pass without an output array for the delta values");
  if (nlag == 0) return PYN_OK;
  cudaStream_t st = pyn::as_stream(stream);
  int64_t chunk = 0;
  const int nchunks = pyn_ibm_chunks(ncand, &chunk);
  const int64_t nwarps = nlag * nchunks;
  PYN_REQUIRE(nwarps * 32 / 256 < 0x7fffffff, "too many (marker, chunk) pairs");
  const unsigned grid = (unsigned)((nwarps * 32 + 255) / 256);
  if (d_ptr)
    ibm_collect_kernel<true><<<grid, 256, 0, st>>>(nlag, d_lagxy, ncand, d_cand, d_nodexy, dl, h, mode, delta_kind,
                                                   nchunks, chunk, d_ptr, d_count_or_nodes, d_dirac);
  else
    ibm_collect_kernel<false><<<grid, 256, 0, st>>>(nlag, d_lagxy, ncand, d_cand, d_nodexy, dl, h, mode, delta_kind,
                                                    nchunks, chunk, nullptr, d_count_or_nodes, nullptr);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_ibm_gram(int64_t nlag, const int64_t* d_ptr, const int32_t* d_nodes, const double* d_dirac, double a, double b,
                 const int64_t* d_aptr, int32_t* d_count_or_cols, double* d_avals, void* stream) {
  PYN_REQUIRE(nlag >= 0 && d_ptr && d_nodes && d_dirac && d_count_or_cols, "bad arguments");
  PYN_REQUIRE(!d_aptr || d_avals, "fill pass without an output array for the values");
  if (nlag == 0) return PYN_OK;
  cudaStream_t st = pyn::as_stream(stream);
  const int nblk = (int)((nlag + 31) / 32);
  const unsigned grid = (unsigned)((nlag * nblk * 32 + 255) / 256);
  if (d_aptr)
    ibm_gram_kernel<true><<<grid, 256, 0, st>>>(nlag, d_ptr, d_nodes, d_dirac, a, b, nblk, d_aptr, d_count_or_cols,
                                                d_avals);
  else
    ibm_gram_kernel<false><<<grid, 256, 0, st>>>(nlag, d_ptr, d_nodes, d_dirac, a, b, nblk, nullptr, d_count_or_cols,
                                                 nullptr);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

}  // extern "C"
