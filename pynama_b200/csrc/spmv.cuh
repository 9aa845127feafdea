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
