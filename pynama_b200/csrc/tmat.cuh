// Template operator: y = A x for a matrix assembled from ONE element matrix on a repetitive mesh, without streaming
// its values.  Interior nodes multiply with the row template of their signature (topo.cuh) through the node adjacency
// -- 4 bytes of HBM per adjacency slot instead of 12 bytes per stored entry --, boundary-affected nodes (the bcmap
// list) use their rows of the assembled CSR.  Same matrix, same result up to summation order.
#pragma once
#include "spmv.cuh"
#include "topo.cuh"

struct pyn_tmat {
  const pyn_topo* topo;      // borrowed
  int rdof, cdof;
  double* T;                 // [sig_slots][rdof][cdof]: per adjacency slot the block of the template (owned)
  const uint8_t* flags;      // borrowed from the bcmap: 1 = row comes from the CSR
  const int* list;           // flagged nodes
  int64_t nlist;
  const double* rowscale;    // borrowed, nullable
  const int64_t* ptr;        // assembled CSR (borrowed): used for the flagged rows only
  const int* ind;
  const double* val;
  int grid_int, grid_bnd;    // persistent grids of the two kernels
};

namespace pyn {

struct TmatView {
  int64_t nnodes;
  const int64_t* adj_ptr;
  const int* adj_cols;
  const int* sig_id;
  const int64_t* sig_ptr;
  const double* T;
  const uint8_t* flags;
  const int* list;
  int64_t nlist;
  const double* rowscale;
  const int64_t* ptr;
  const int* ind;
  const double* val;
};
static inline TmatView view_of(const pyn_tmat* m) {
  TmatView v;
  v.nnodes = m->topo->nnodes; v.adj_ptr = m->topo->adj_ptr; v.adj_cols = m->topo->adj_cols;
  v.sig_id = m->topo->sig_id; v.sig_ptr = m->topo->sig_ptr; v.T = m->T; v.flags = m->flags; v.list = m->list;
  v.nlist = m->nlist; v.rowscale = m->rowscale; v.ptr = m->ptr; v.ind = m->ind; v.val = m->val;
  return v;
}

// One warp computes the RD rows of interior node r; the sums are valid in lane 0.
template <int RD, int CD>
__device__ __forceinline__ void tmat_node(const TmatView& m, int64_t r, int lane, const double* __restrict__ x,
                                          double (&acc)[RD]) {
  const int64_t a0 = m.adj_ptr[r];
  const int L = (int)(m.adj_ptr[r + 1] - a0);
  const double* Ts = m.T + m.sig_ptr[m.sig_id[r]] * (RD * CD);
  const double rs = m.rowscale ? m.rowscale[r] : 1.0;
#pragma unroll
  for (int a = 0; a < RD; ++a) acc[a] = 0.0;
  for (int p = lane; p < L; p += 32) {
    const int c = ld_stream(m.adj_cols + a0 + p);
    double xv[CD];
    if (CD == 2) {
      const double2 t = *reinterpret_cast<const double2*>(x + (int64_t)c * 2);
      xv[0] = t.x;
      xv[CD - 1] = t.y;
    } else {
#pragma unroll
      for (int b = 0; b < CD; ++b) xv[b] = x[(int64_t)c * CD + b];
    }
    const double* Tb = Ts + p * (RD * CD);
#pragma unroll
    for (int a = 0; a < RD; ++a)
#pragma unroll
      for (int b = 0; b < CD; ++b) {
        double tv = __ldg(Tb + a * CD + b);
        if (m.rowscale) tv = rs * tv;          // the assembled entry is rowscale * template
        acc[a] += tv * xv[b];
      }
  }
#pragma unroll
  for (int a = 0; a < RD; ++a) acc[a] = warp_sum(acc[a]);
}

}  // namespace pyn
