// Element operator ("template operator" of the C ABI, pyn_tmat_*): y = A x for a matrix assembled from ONE element
// matrix E on a repetitive mesh, WITHOUT streaming its stored values.
//
//   phase 1 (eop_apply_kernel, tmat.cu): Ye[e][:] = E . x[dofs(e)] for every element -- a batched dense product
//            [NR x NC] . [NC x ne] with E resident in shared memory, 128 elements per tile, each thread an R-row x
//            4-element register block (R = 8 | 10); HBM traffic = x once + 4 B/element-node of connectivity + Ye written once.
//   phase 2 (eop_gather_node, below): row node r sums Ye[e][loc_e(r)] over its incident elements in ascending cell
//            order (the reference's cell loop order, base_problem.py:499-552), scaled by rowscale[r] -- a deterministic
//            gather, no atomics.
//   boundary-affected nodes (bcmap list) take their rows from the assembled CSR (one warp per row).
//   Layout of Ye (rowmap): the rows of an element are split by the KIND of their node -- the 4 vertices, each of the 4
//   edges, the cell interior (the reference's element-local order, indices.py:66-88) -- and every kind is its own dense
//   array over the elements.  Phase 2 walks the nodes in global order, which is kind by kind (cell interiors, vertices,
//   edges: dmplex.py:42-61), so each of its passes reads one dense array instead of 16-byte pieces scattered in
//   element-major records (DRAM sectors fully used: 1.85 -> ~1.1 GB moved per product at cfg 4).
//
// Algorithmic HBM bytes per product (K at 1024^2 x ngl 5: 1.6 GB against 29.8 GB of the CSR stream):
//   8 cols + 4 ne nn + 2 * 8 ne NR + 8 (nnodes+1) + 6 ne nn + 8 rows.
// Same operator, same result up to summation order (sum over elements of row sums instead of row sum of sums).
#pragma once
#include "spmv.cuh"
#include "topo.cuh"

struct pyn_tmat {
  const pyn_topo* topo;      // borrowed
  int rdof, cdof;
  int NR, NC, NRP, NCP;      // element matrix rows / cols; rows padded to a multiple of R, cols to 2*odd
  int R, W;                  // rows per register block (8 | 10), warps per CTA of phase 1
  int ag;                    // DMMA variant with the A fragments read through L1/L2 instead of shared memory (ngl 8-9)
  int mma, RB, KB, MB, P;    // DMMA variant of phase 1: 8-row blocks, 4-column blocks (Et then holds the fragments),
                             // row blocks per pass, gathered pairs per lane and stage
  double* Et;                // owned: sign * E transposed, [NCP][NRP], zero padded
  double* Ye;                // owned scratch: NR values per element, laid out by rowmap
  int2* rowmap;              // owned [NR]: Ye index of (element e, row i) = rowmap[i].x + e * rowmap[i].y
  const uint8_t* flags;      // borrowed from the bcmap: 1 = row comes from the CSR
  const int* list;           // flagged nodes
  int64_t nlist;
  const double* rowscale;    // borrowed, nullable
  const int64_t* ptr;        // assembled CSR (borrowed): used for the flagged rows only
  const int* ind;
  const double* val;
  size_t smem;               // dynamic shared memory of eop_apply_kernel
  int grid_apply, grid_int, grid_bnd;   // persistent grids: phase 1, phase 2, flagged rows
};

namespace pyn {

constexpr int EO_TE = 128;       // elements per tile
constexpr int EO_Q = 4;          // elements per register block (per lane)

struct TmatView {
  int64_t nnodes, ne;
  int nn, NR, NC, NRP, NCP;
  unsigned nn_magic;   // ceil(2^32 / nn): i / nn == __umulhi(i, nn_magic) for i * nn < 2^32
  const int* e2n;
  const int64_t* n2e_ptr;
  const int* n2e_elem;
  const short* n2e_loc;
  const double* Et;
  double* Ye;
  const int2* rowmap;
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
  v.nnodes = m->topo->nnodes; v.ne = m->topo->ne; v.nn = m->topo->nn; v.NR = m->NR; v.NC = m->NC; v.NRP = m->NRP; v.NCP = m->NCP;
  v.nn_magic = (unsigned)((0x100000000ull + (unsigned)v.nn - 1) / (unsigned)v.nn);
  v.e2n = m->topo->e2n; v.n2e_ptr = m->topo->n2e_ptr; v.n2e_elem = m->topo->n2e_elem; v.n2e_loc = m->topo->n2e_loc;
  v.Et = m->Et; v.Ye = m->Ye; v.rowmap = m->rowmap; v.flags = m->flags; v.list = m->list; v.nlist = m->nlist; v.rowscale = m->rowscale;
  v.ptr = m->ptr; v.ind = m->ind; v.val = m->val;
  return v;
}

// phase 1 on `st` (defined in tmat.cu); `done` (nullable): device flag, non-zero = skip (CG convergence flag)
int tmat_apply(const pyn_tmat* m, const double* d_x, const int* d_done, cudaStream_t st);
// allocations and launch shapes of a handle whose topo / rdof / cdof / bcmap / CSR fields are set (pyn_tmat_create)
int tmat_init(pyn_tmat* m, const double* d_E, double sign, cudaStream_t st);

// phase 2 for one un-flagged node: acc[a] = rowscale[r] * sum over incident elements (ascending) of Ye[e][loc*RD + a]
template <int RD>
__device__ __forceinline__ void eop_gather_node(const TmatView& m, int64_t r, double (&acc)[RD]) {
  const int64_t b0 = m.n2e_ptr[r], b1 = m.n2e_ptr[r + 1];
#pragma unroll
  for (int a = 0; a < RD; ++a) acc[a] = 0.0;
  for (int64_t k = b0; k < b1; ++k) {
    const int e = ld_stream(m.n2e_elem + k);
    const int loc = m.n2e_loc[k];
    const int2 rm = __ldg(m.rowmap + loc * RD);
    const double* p = m.Ye + rm.x + (int64_t)e * rm.y;
    if (RD == 2) {
      const double2 t = *reinterpret_cast<const double2*>(p);   // bases and strides of the row map are even
      acc[0] += t.x;
      acc[RD - 1] += t.y;
    } else {
#pragma unroll
      for (int a = 0; a < RD; ++a) acc[a] += p[a];
    }
  }
  if (m.rowscale) {
    const double rs = m.rowscale[r];
#pragma unroll
    for (int a = 0; a < RD; ++a) acc[a] = rs * acc[a];
  }
}

}  // namespace pyn
