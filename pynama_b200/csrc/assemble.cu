// Deterministic global assembly into CSR.
//
// Row-gather formulation (no atomics): one warp owns one mesh node = `rdof` consecutive matrix rows.  It walks the
// node's elements in ASCENDING cell order and adds each element block into shared-memory accumulators indexed by
// the node's sorted adjacency list, then writes the rows out compacted by the boundary-condition masks.  The
// summation order is therefore exactly the reference's serial cell loop (base_problem.py:337,504,135) and the
// result is bitwise reproducible run to run.
#include <cub/cub.cuh>

#include "common.cuh"
#include "topo.cuh"

namespace {

struct DescView {
  int rdof, cdof;
  const int8_t* rclass;
  const int8_t* cclass;
  uint32_t pair_mask, diag_classes;
  double diag_add[4];
  double sign;
  const double* E;
  int64_t E_stride;
  const double* rowscale;
};

__device__ __forceinline__ int lower_bound_smem(const int* __restrict__ a, int n, int key) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (a[mid] < key) lo = mid + 1; else hi = mid;
  }
  return lo;
}

template <bool COUNT>
__global__ void __launch_bounds__(128)
asm_rows_kernel(pyn::TopoView t, DescView d, int Lmax, int* __restrict__ rowcnt, const int64_t* __restrict__ indptr,
                int* __restrict__ indices, double* __restrict__ data) {
  extern __shared__ double smd[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int rd = d.rdof, cd = d.cdof, blk = rd * cd;
  const int per_warp_dbl = COUNT ? (Lmax + 1) / 2 : Lmax * blk + (Lmax + 1) / 2;
  double* acc = smd + (size_t)w * per_warp_dbl;
  int* scols = COUNT ? (int*)acc : (int*)(acc + Lmax * blk);
  const bool values = (!COUNT) && (data != nullptr);
  const bool square = (rd == cd);

  for (int64_t r = blockIdx.x * (int64_t)wpb + w; r < t.nnodes; r += (int64_t)gridDim.x * wpb) {
    const int64_t a0 = t.adj_ptr[r];
    const int L = (int)(t.adj_ptr[r + 1] - a0);
    for (int i = lane; i < L; i += 32) scols[i] = t.adj_cols[a0 + i];
    if (values)
      for (int i = lane; i < L * blk; i += 32) acc[i] = 0.0;
    __syncwarp();
    if (values) {
      const int64_t b0 = t.n2e_ptr[r];
      const int deg = (int)(t.n2e_ptr[r + 1] - b0);
      for (int k = 0; k < deg; ++k) {  // ascending cell order
        const int e = t.n2e_elem[b0 + k];
        const int lr = t.n2e_loc[b0 + k];
        const double* Ee = d.E + (d.E_stride ? (int64_t)e * d.E_stride : 0) + (int64_t)lr * rd * (cd * t.nn);
        const int* en = t.e2n + (int64_t)e * t.nn;
        for (int l = lane; l < t.nn; l += 32) {
          const int c = en[l];
          const int p = lower_bound_smem(scols, L, c);
          for (int a = 0; a < rd; ++a)
            for (int b = 0; b < cd; ++b)
              acc[(p * rd + a) * cd + b] += d.sign * Ee[(int64_t)a * (cd * t.nn) + l * cd + b];
        }
        __syncwarp();
      }
    }
    const double rs = (values && d.rowscale) ? d.rowscale[r] : 1.0;
    for (int a = 0; a < rd; ++a) {
      const int64_t row = r * rd + a;
      const int rca = d.rclass ? d.rclass[row] : 0;
      const bool dgrow = square && ((d.diag_classes >> rca) & 1u);
      const int64_t base = COUNT ? 0 : indptr[row];
      int running = 0;
      const int tot = L * cd;
      for (int t0 = 0; t0 < tot; t0 += 32) {
        const int tt = t0 + lane;
        const bool valid = tt < tot;
        const int p = valid ? tt / cd : 0, b = valid ? tt % cd : 0;
        const int c = scols[p];
        const int cc = d.cclass ? d.cclass[(int64_t)c * cd + b] : 0;
        const bool pa = (d.pair_mask >> (rca * 4 + cc)) & 1u;
        const bool dg = dgrow && (c == r) && (a == b);
        const bool keep = valid && (pa || dg);
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (!COUNT && keep) {
          const int64_t pos = base + running + __popc(m & ((1u << lane) - 1));
          if (indices) indices[pos] = c * cd + b;
          if (values) {
            double v = pa ? acc[(p * rd + a) * cd + b] : 0.0;
            if (d.rowscale) v = rs * v;
            if (dg) v += d.diag_add[rca];
            data[pos] = v;
          }
        }
        running += __popc(m);
      }
      if (COUNT && lane == 0) rowcnt[row] = running;
    }
    __syncwarp();
  }
}

__global__ void weights_kernel(pyn::TopoView t, const double* __restrict__ We, int64_t stride, double* __restrict__ W) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < t.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    double s = 0.0;
    for (int64_t k = t.n2e_ptr[r]; k < t.n2e_ptr[r + 1]; ++k)
      s += We[(stride ? (int64_t)t.n2e_elem[k] * stride : 0) + t.n2e_loc[k]];
    W[r] = s;
  }
}

// ---- CSR utilities -------------------------------------------------------------------------------------------
template <bool COUNT>
__global__ void csr_add_kernel(int64_t nrows, const int64_t* __restrict__ ap, const int* __restrict__ ai,
                               const double* __restrict__ av, double alpha, const int64_t* __restrict__ bp,
                               const int* __restrict__ bi, const double* __restrict__ bv, int* __restrict__ cnt,
                               const int64_t* __restrict__ cp, int* __restrict__ ci, double* __restrict__ cv) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
    int64_t i = ap[r], ie = ap[r + 1], j = bp[r], je = bp[r + 1];
    int64_t o = COUNT ? 0 : cp[r];
    int n = 0;
    while (i < ie || j < je) {
      const int ca = (i < ie) ? ai[i] : 0x7fffffff, cb = (j < je) ? bi[j] : 0x7fffffff;
      if (!COUNT) {
        // MatAXPY: y_ij + alpha * x_ij ; entries present in only one operand keep that operand's value
        double v;
        if (ca == cb) v = av[i] + alpha * bv[j];
        else if (ca < cb) v = av[i];
        else v = alpha * bv[j];
        ci[o] = ca < cb ? ca : cb;
        cv[o] = v;
        ++o;
      }
      ++n;
      if (ca <= cb) ++i;
      if (cb <= ca) ++j;
    }
    if (COUNT) cnt[r] = n;
  }
}

__global__ void csr_diag_kernel(int64_t nrows, const int64_t* __restrict__ p, const int* __restrict__ ind,
                                const double* __restrict__ val, double* __restrict__ diag) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
    double dv = 0.0;
    int64_t lo = p[r], hi = p[r + 1];
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (ind[mid] < r) lo = mid + 1; else hi = mid;
    }
    if (lo < p[r + 1] && ind[lo] == r) dv = val[lo];
    diag[r] = dv;
  }
}

template <bool COUNT>
__global__ void csr_compress_kernel(int64_t nrows, const int64_t* __restrict__ p, const int* __restrict__ ind,
                                    const double* __restrict__ val, int* __restrict__ cnt,
                                    const int64_t* __restrict__ op, int* __restrict__ oind, double* __restrict__ oval) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp0; r < nrows; r += nw) {
    const int64_t b = p[r], e = p[r + 1];
    const int64_t o = COUNT ? 0 : op[r];
    int running = 0;
    for (int64_t i0 = b; i0 < e; i0 += 32) {
      const int64_t i = i0 + lane;
      const double v = (i < e) ? val[i] : 0.0;
      const bool keep = (i < e) && (v != 0.0);
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (!COUNT && keep) {
        const int64_t pos = o + running + __popc(m & ((1u << lane) - 1));
        oind[pos] = ind[i];
        oval[pos] = v;
      }
      running += __popc(m);
    }
    if (COUNT && lane == 0) cnt[r] = running;
  }
}

// ---- generic COO -> CSR ----------------------------------------------------------------------------------------
struct CooScratch {
  int64_t n, nrows, ncols, nnz;
  int64_t* keys;   // sorted keys
  int64_t* perm;   // original position of each sorted entry
  int64_t* seg;    // unique-entry id of each sorted entry
};

__global__ void coo_keys_kernel(int64_t n, const int* __restrict__ rows, const int* __restrict__ cols, int64_t ncols,
                                int64_t* __restrict__ keys, int64_t* __restrict__ idx) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    keys[i] = (int64_t)rows[i] * ncols + cols[i];
    idx[i] = i;
  }
}
__global__ void coo_heads_kernel(int64_t n, const int64_t* __restrict__ keys, int64_t ncols, int64_t* __restrict__ head,
                                 int* __restrict__ rowcnt) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const bool h = (i == 0) || (keys[i] != keys[i - 1]);
    head[i] = h ? 1 : 0;
    if (h) atomicAdd(&rowcnt[keys[i] / ncols], 1);
  }
}
__global__ void coo_fill_kernel(int64_t n, const int64_t* __restrict__ keys, const int64_t* __restrict__ perm,
                                const int64_t* __restrict__ seg, int64_t ncols, const double* __restrict__ vals,
                                int* __restrict__ indices, double* __restrict__ data) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    if (i != 0 && keys[i] == keys[i - 1]) continue;  // only segment heads work
    const int64_t k = keys[i];
    const int64_t o = seg[i] - 1;
    if (indices) indices[o] = (int)(k % ncols);
    if (data) {
      double s = 0.0;
      for (int64_t j = i; j < n && keys[j] == k; ++j) s += vals[perm[j]];  // input (= cell) order: the sort is stable
      data[o] = s;
    }
  }
}

int make_desc(const pyn_asm_desc* desc, DescView* v) {
  PYN_REQUIRE(desc, "desc is NULL");
  PYN_REQUIRE(desc->rdof >= 1 && desc->rdof <= 3 && desc->cdof >= 1 && desc->cdof <= 3, "rdof/cdof must be in 1..3");
  v->rdof = desc->rdof; v->cdof = desc->cdof;
  v->rclass = desc->d_rclass; v->cclass = desc->d_cclass;
  v->pair_mask = desc->pair_mask; v->diag_classes = desc->diag_classes;
  for (int i = 0; i < 4; ++i) v->diag_add[i] = desc->diag_add[i];
  v->sign = desc->sign; v->E = desc->d_E; v->E_stride = desc->E_stride; v->rowscale = desc->d_rowscale;
  return PYN_OK;
}

}  // namespace

extern "C" {

int pyn_asm_count(const pyn_topo* t, const pyn_asm_desc* desc, int64_t* d_indptr, int64_t* h_nnz, void* stream) {
  PYN_REQUIRE(t && d_indptr, "bad arguments");
  DescView dv;
  int rc = make_desc(desc, &dv);
  if (rc) return rc;
  cudaStream_t st = pyn::as_stream(stream);
  const int64_t nrows = t->nnodes * dv.rdof;
  int* cnt = nullptr;
  PYN_CHECK_CUDA(cudaMallocAsync(&cnt, sizeof(int) * (nrows + 1), st));
  const int wpb = 4;
  const size_t smem = (size_t)wpb * ((t->max_row + 1) / 2) * sizeof(double);
  asm_rows_kernel<true><<<pyn::grid_for(t->nnodes, wpb, 16), wpb * 32, smem, st>>>(pyn::view_of(t), dv, t->max_row, cnt,
                                                                                   nullptr, nullptr, nullptr);
  PYN_CHECK_LAUNCH();
  rc = pyn::scan_counts(cnt, nrows, d_indptr, h_nnz, st);
  PYN_CHECK_CUDA(cudaFreeAsync(cnt, st));
  return rc;
}

int pyn_asm_fill(const pyn_topo* t, const pyn_asm_desc* desc, const int64_t* d_indptr, int32_t* d_indices,
                 double* d_data, void* stream) {
  PYN_REQUIRE(t && d_indptr, "bad arguments");
  DescView dv;
  int rc = make_desc(desc, &dv);
  if (rc) return rc;
  PYN_REQUIRE(d_data == nullptr || dv.E != nullptr, "element matrices missing");
  const int wpb = 4;
  const int blk = dv.rdof * dv.cdof;
  const size_t smem = (size_t)wpb * ((size_t)t->max_row * blk + (t->max_row + 1) / 2) * sizeof(double);
  if (smem > 227 * 1024) {
    pyn::set_error("pyn_asm_fill: adjacency row length %d needs %zu B shared memory", t->max_row, smem);
    return PYN_ERR_LIMIT;
  }
  PYN_CHECK_CUDA(cudaFuncSetAttribute(asm_rows_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  asm_rows_kernel<false><<<pyn::grid_for(t->nnodes, wpb, 16), wpb * 32, smem, pyn::as_stream(stream)>>>(
      pyn::view_of(t), dv, t->max_row, nullptr, d_indptr, d_indices, d_data);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_asm_weights(const pyn_topo* t, const double* d_We, int64_t stride, double* d_W, void* stream) {
  PYN_REQUIRE(t && d_We && d_W, "bad arguments");
  weights_kernel<<<pyn::grid_for(t->nnodes, 256), 256, 0, pyn::as_stream(stream)>>>(pyn::view_of(t), d_We, stride, d_W);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_csr_add_count(int64_t nrows, const int64_t* ap, const int32_t* ai, const int64_t* bp, const int32_t* bi,
                      int64_t* cp, int64_t* h_nnz, void* stream) {
  PYN_REQUIRE(nrows > 0 && ap && ai && bp && bi && cp, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  int* cnt = nullptr;
  PYN_CHECK_CUDA(cudaMallocAsync(&cnt, sizeof(int) * (nrows + 1), st));
  csr_add_kernel<true><<<pyn::grid_for(nrows, 128), 128, 0, st>>>(nrows, ap, ai, nullptr, 0.0, bp, bi, nullptr, cnt,
                                                                  nullptr, nullptr, nullptr);
  PYN_CHECK_LAUNCH();
  int rc = pyn::scan_counts(cnt, nrows, cp, h_nnz, st);
  PYN_CHECK_CUDA(cudaFreeAsync(cnt, st));
  return rc;
}

int pyn_csr_add_fill(int64_t nrows, const int64_t* ap, const int32_t* ai, const double* av, double alpha,
                     const int64_t* bp, const int32_t* bi, const double* bv, const int64_t* cp, int32_t* ci, double* cv,
                     void* stream) {
  PYN_REQUIRE(nrows > 0 && ap && ai && av && bp && bi && bv && cp && ci && cv, "bad arguments");
  csr_add_kernel<false><<<pyn::grid_for(nrows, 128), 128, 0, pyn::as_stream(stream)>>>(nrows, ap, ai, av, alpha, bp, bi,
                                                                                        bv, nullptr, cp, ci, cv);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_csr_diagonal(int64_t nrows, const int64_t* p, const int32_t* ind, const double* val, double* diag, void* stream) {
  PYN_REQUIRE(nrows > 0 && p && ind && val && diag, "bad arguments");
  csr_diag_kernel<<<pyn::grid_for(nrows, 256), 256, 0, pyn::as_stream(stream)>>>(nrows, p, ind, val, diag);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_csr_compress_count(int64_t nrows, const int64_t* p, const double* val, int64_t* op, int64_t* h_nnz, void* stream) {
  PYN_REQUIRE(nrows > 0 && p && val && op, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  int* cnt = nullptr;
  PYN_CHECK_CUDA(cudaMallocAsync(&cnt, sizeof(int) * (nrows + 1), st));
  csr_compress_kernel<true><<<pyn::grid_for(nrows * 32, 256), 256, 0, st>>>(nrows, p, nullptr, val, cnt, nullptr, nullptr,
                                                                            nullptr);
  PYN_CHECK_LAUNCH();
  int rc = pyn::scan_counts(cnt, nrows, op, h_nnz, st);
  PYN_CHECK_CUDA(cudaFreeAsync(cnt, st));
  return rc;
}

int pyn_csr_compress_fill(int64_t nrows, const int64_t* p, const int32_t* ind, const double* val, const int64_t* op,
                          int32_t* oind, double* oval, void* stream) {
  PYN_REQUIRE(nrows > 0 && p && ind && val && op && oind && oval, "bad arguments");
  csr_compress_kernel<false><<<pyn::grid_for(nrows * 32, 256), 256, 0, pyn::as_stream(stream)>>>(nrows, p, ind, val,
                                                                                                 nullptr, op, oind, oval);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_coo_count(int64_t n, const int32_t* d_rows, const int32_t* d_cols, int64_t nrows, int64_t ncols,
                  int64_t* d_indptr, int64_t* h_nnz, void** scratch, void* stream) {
  PYN_REQUIRE(n > 0 && d_rows && d_cols && nrows > 0 && ncols > 0 && d_indptr && scratch, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  CooScratch* s = new CooScratch();
  memset(s, 0, sizeof(*s));
  s->n = n; s->nrows = nrows; s->ncols = ncols;
  int64_t *k0 = nullptr, *i0 = nullptr;
  PYN_CHECK_CUDA(cudaMalloc(&k0, sizeof(int64_t) * n));
  PYN_CHECK_CUDA(cudaMalloc(&i0, sizeof(int64_t) * n));
  PYN_CHECK_CUDA(cudaMalloc(&s->keys, sizeof(int64_t) * n));
  PYN_CHECK_CUDA(cudaMalloc(&s->perm, sizeof(int64_t) * n));
  PYN_CHECK_CUDA(cudaMalloc(&s->seg, sizeof(int64_t) * n));
  coo_keys_kernel<<<pyn::grid_for(n, 256), 256, 0, st>>>(n, d_rows, d_cols, ncols, k0, i0);
  PYN_CHECK_LAUNCH();
  int end_bit = 1;
  while (end_bit < 63 && ((int64_t)1 << end_bit) < nrows * ncols) ++end_bit;
  void* tmp = nullptr; size_t bytes = 0;
  PYN_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, k0, s->keys, i0, s->perm, n, 0, end_bit, st));
  PYN_CHECK_CUDA(cudaMalloc(&tmp, bytes));
  PYN_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(tmp, bytes, k0, s->keys, i0, s->perm, n, 0, end_bit, st));  // stable
  pyn::count_launch(8);
  int* cnt = nullptr;
  PYN_CHECK_CUDA(cudaMalloc(&cnt, sizeof(int) * (nrows + 1)));
  PYN_CHECK_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int) * (nrows + 1), st));
  coo_heads_kernel<<<pyn::grid_for(n, 256), 256, 0, st>>>(n, s->keys, ncols, k0 /* reuse as head flags */, cnt);
  PYN_CHECK_LAUNCH();
  void* tmp2 = nullptr; size_t bytes2 = 0;
  PYN_CHECK_CUDA(cub::DeviceScan::InclusiveSum(tmp2, bytes2, k0, s->seg, n, st));
  PYN_CHECK_CUDA(cudaMalloc(&tmp2, bytes2));
  PYN_CHECK_CUDA(cub::DeviceScan::InclusiveSum(tmp2, bytes2, k0, s->seg, n, st));
  pyn::count_launch(2);
  int rc = pyn::scan_counts(cnt, nrows, d_indptr, &s->nnz, st);
  cudaFree(tmp); cudaFree(tmp2); cudaFree(k0); cudaFree(i0); cudaFree(cnt);
  if (rc) return rc;
  if (h_nnz) *h_nnz = s->nnz;
  *scratch = s;
  return PYN_OK;
}

int pyn_coo_fill(void* scratch, const double* d_vals, const int64_t* d_indptr, int32_t* d_indices, double* d_data,
                 void* stream) {
  PYN_REQUIRE(scratch, "scratch is NULL");
  CooScratch* s = (CooScratch*)scratch;
  (void)d_indptr;
  cudaStream_t st = pyn::as_stream(stream);
  PYN_REQUIRE(d_data == nullptr || d_vals != nullptr, "values missing");
  coo_fill_kernel<<<pyn::grid_for(s->n, 256), 256, 0, st>>>(s->n, s->keys, s->perm, s->seg, s->ncols, d_vals, d_indices,
                                                            d_data);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  cudaFree(s->keys); cudaFree(s->perm); cudaFree(s->seg);
  delete s;
  return PYN_OK;
}
}
