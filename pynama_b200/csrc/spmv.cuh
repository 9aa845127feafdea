// CSR SpMV building block shared by spmv.cu and ksp.cu.
//
// "CSR-vector": a sub-warp of LPR lanes owns one row; lanes stride through the row so that the value / index
// streams are read with coalesced, L1-bypassing loads, the gathered vector goes through the normal cached path,
// and the row sum is finished with a fixed shuffle tree (deterministic).  Rows are distributed grid-stride so the
// number of CTAs (and of per-CTA dot-product partials) is bounded by the SM count.
#pragma once
#include "common.cuh"

namespace pyn {

template <int LPR>
__device__ __forceinline__ double spmv_row(const int64_t* __restrict__ ptr, const int* __restrict__ ind,
                                           const double* __restrict__ val, const double* __restrict__ x, int64_t row,
                                           bool active, int sub) {
  double s0 = 0.0, s1 = 0.0;
  if (active) {
    const int64_t b = ptr[row], e = ptr[row + 1];
    int64_t i = b + sub;
    for (; i + LPR < e; i += 2 * LPR) {  // two independent loads in flight per lane
      const double v0 = ld_stream(val + i), v1 = ld_stream(val + i + LPR);
      const int c0 = ld_stream(ind + i), c1 = ld_stream(ind + i + LPR);
      s0 += v0 * x[c0];
      s1 += v1 * x[c1];
    }
    if (i < e) s0 += ld_stream(val + i) * x[ld_stream(ind + i)];
  }
  return subwarp_sum<LPR>(s0 + s1);
}

static inline int pick_lpr(int64_t nrows, int64_t nnz) {
  const double avg = nrows > 0 ? (double)nnz / (double)nrows : 0.0;
  if (avg <= 6.0) return 4;
  if (avg <= 12.0) return 8;
  if (avg <= 40.0) return 16;
  return 32;
}

}  // namespace pyn

// ---------------------------------------------------------------------------------------------------------------
// "CSR-stream" SpMV: the matrix is cut into chunks of ~SP_C consecutive non-zeros snapped to row boundaries (plan
// built once per matrix).  A CTA streams one chunk's values and column indices with fully coalesced, independent
// loads (8 in flight per thread), multiplies by the gathered x and parks the products in shared memory; the row
// sums are then taken from shared memory by sub-warps whose width adapts to the number of rows in the chunk.  HBM
// efficiency therefore does not depend on the row lengths (1 ... several hundred in these matrices), and the
// summation order is a pure function of the sparsity pattern (bitwise reproducible).
// ---------------------------------------------------------------------------------------------------------------
namespace pyn {

constexpr int SP_BLK = 256;      // threads per CTA
constexpr int SP_U = 8;          // independent (value, index) loads in flight per thread
constexpr int SP_CAP = 3072;     // products held in shared memory per chunk
constexpr int SP_RCAP = 512;     // row offsets staged in shared memory per chunk

struct SpmvPlanView {
  int64_t nrows, nnz;
  const int64_t* ptr;
  const int* blk_row;  // [nblk+1] first row of each chunk
  int nblk;
};

struct SpmvSmem {
  double prod[SP_CAP];
  int off[SP_RCAP + 1];
};

// Processes chunk `b`; calls epi(row, sum) once per row from exactly one thread.
template <class Epi>
__device__ __forceinline__ void spmv_stream_chunk(const SpmvPlanView& P, const int* __restrict__ ind,
                                                  const double* __restrict__ val, const double* __restrict__ x, int b,
                                                  SpmvSmem& sm, Epi&& epi) {
  const int tid = threadIdx.x;
  const int r0 = P.blk_row[b], r1 = P.blk_row[b + 1];
  const int nr = r1 - r0;
  if (nr <= 0) return;
  const int64_t p0 = P.ptr[r0];
  const int cnt = (int)min(P.ptr[r1] - p0, (int64_t)0x7fffffff);
  if (cnt <= SP_CAP) {
    const bool staged = nr <= SP_RCAP;
    if (staged)
      for (int i = tid; i <= nr; i += SP_BLK) sm.off[i] = (int)(P.ptr[r0 + i] - p0);
    const double* v0 = val + p0;
    const int* c0 = ind + p0;
    for (int base = 0; base < cnt; base += SP_U * SP_BLK) {
      double vv[SP_U];
      int cc[SP_U];
#pragma unroll
      for (int u = 0; u < SP_U; ++u) {
        const int i = base + u * SP_BLK + tid;
        const bool ok = i < cnt;
        vv[u] = ok ? ld_stream(v0 + i) : 0.0;
        cc[u] = ok ? ld_stream(c0 + i) : -1;
      }
#pragma unroll
      for (int u = 0; u < SP_U; ++u) {
        const int i = base + u * SP_BLK + tid;
        if (cc[u] >= 0) sm.prod[i] = vv[u] * x[cc[u]];
      }
    }
    __syncthreads();
    int lpr = 32;
    while (lpr > 1 && lpr * nr > SP_BLK) lpr >>= 1;
    const int rpp = SP_BLK / lpr, sub = tid & (lpr - 1), slot = tid / lpr;
    for (int rb = 0; rb < nr; rb += rpp) {
      const int i = rb + slot;
      const bool active = i < nr;
      double s = 0.0;
      if (active) {
        const int lo = staged ? sm.off[i] : (int)(P.ptr[r0 + i] - p0);
        const int hi = staged ? sm.off[i + 1] : (int)(P.ptr[r0 + i + 1] - p0);
        for (int j = lo + sub; j < hi; j += lpr) s += sm.prod[j];
      }
      for (int o = lpr >> 1; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o, lpr);
      if (active && sub == 0) epi((int64_t)r0 + i, s);
    }
    __syncthreads();
  } else {
    // a chunk holding a row longer than the shared-memory budget: one warp per row straight from global memory
    const int lane = tid & 31, w = tid >> 5;
    for (int i = w; i < nr; i += SP_BLK / 32) {
      const double s = spmv_row<32>(P.ptr, ind, val, x, (int64_t)r0 + i, true, lane);
      if (lane == 0) epi((int64_t)r0 + i, s);
    }
  }
}

}  // namespace pyn

struct pyn_spmv_plan {
  int64_t nrows, nnz;
  const int64_t* ptr;  // borrowed
  int* blk_row;
  int nblk, chunk, grid;
};
namespace pyn {
static inline SpmvPlanView view_of(const pyn_spmv_plan* p) {
  SpmvPlanView v;
  v.nrows = p->nrows; v.nnz = p->nnz; v.ptr = p->ptr; v.blk_row = p->blk_row; v.nblk = p->nblk;
  return v;
}
}  // namespace pyn
