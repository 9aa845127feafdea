/* pynama_b200 -- C ABI of the B200-native FEM hot path (libpynama_b200.so)
 *
 * The reference (ibejarano/Pynama) is pure Python on top of petsc4py and defines no FFI of its own; the native
 * code it executes on this path is PETSc's.  Each entry point below therefore cites the reference call site
 * (path:line under /root/reference/) whose work it replaces; INTEGRATION.md shows the ctypes binding a
 * maintainer would add on the reference side.
 *
 * Conventions
 *   - plain C: raw pointers and sizes only.  Every `d_` pointer is a DEVICE pointer owned by the caller;
 *     `h_` pointers are host pointers.  No hidden allocation except inside opaque handles
 *     (`*_create` / `*_destroy`).
 *   - every function returns 0 on success, a negative pyn_status on failure; pyn_last_error() gives the text.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Calls are asynchronous with
 *     respect to the host unless stated otherwise.
 *   - all values FP64, column indices int32, row pointers int64, dof classes int8.
 *   - vector layout: dof = node*ncomp + comp (reference: indices.py:90-92).
 *   - one host thread per GPU; nothing here is re-entrant on the same handle.
 */
#ifndef PYNAMA_B200_H
#define PYNAMA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  PYN_OK = 0,
  PYN_ERR_ARG = -1,      /* bad argument */
  PYN_ERR_CUDA = -2,     /* CUDA runtime error (text in pyn_last_error) */
  PYN_ERR_LIMIT = -3,    /* size outside the compiled limits (ngl, node valence) */
  PYN_ERR_NCCL = -4,
  PYN_ERR_BREAKDOWN = -5 /* Krylov breakdown / indefinite operator */
} pyn_status;

#define PYN_MAX_NGL 12

const char* pyn_last_error(void);
int pyn_version(void);
/* number of kernel launches issued by this library on this process since load (bench.py: gpu_launches) */
int64_t pyn_launch_count(void);
int pyn_device_info(int* sm_count, int* cc_major, int* cc_minor, int64_t* smem_per_block_optin);

/* ------------------------------------------------------------------------------------------------------
 * Element layer.  Replaces Spectral.getElemKLEMatrices / getElemKLEOperators
 * (src/elements/spectral.py:89-157, :159-218) for dim = 2, batched over elements.
 * The 1-D tables are the host-computed Q1-Q3 tables (src/elements/utilities.py:43-92,
 * src/elements/element.py:17-48):  h[q*ngl+a] = l_a(xi_q), d[q*ngl+a] = l_a'(xi_q).
 * perm[k] = a*ngl+b : element-local node k -> tensor index (src/elements/spectral.py:224-246).
 * ---------------------------------------------------------------------------------------------------- */
typedef struct pyn_elem pyn_elem;

int pyn_elem_create(int ngl,
                    int nq_full, const double* h_full, const double* d_full, const double* w_full, const double* xi_full,
                    int nq_red, const double* h_red, const double* d_red, const double* w_red, const double* xi_red,
                    const double* d_nodal /* [ngl*ngl] l_a'(node_q) */, const double* w_nodal, const double* xi_nodal,
                    const int32_t* perm /* [ngl*ngl] */, double alpha_w, double alpha_d, pyn_elem** out);
int pyn_elem_destroy(pyn_elem* e);

/* d_coords [ne][4][2] corner coordinates in closure order.  Outputs (any may be NULL to skip):
 * d_K [ne][2nn][2nn], d_Rw [ne][2nn][nn], d_Rd [ne][2nn][nn]; index = localNode*2 + comp. */
int pyn_elem_kle(const pyn_elem* e, const double* d_coords, int64_t ne,
                 double* d_K, double* d_Rw, double* d_Rd, void* stream);
/* Affine elements (parallelograms: every box mesh of the reference): the same quadrature in closed form -- 17
 * coefficients per element and six 1-D matrices (csrc/elem_affine.cuh, SURVEY 8a').  pyn_elem_affine_coef writes
 * d_coef [ne][18] and returns the number of NON-affine elements (synchronises when h_nonaffine is given): zero means the
 * mesh can take pyn_elem_kle_affine (materialised element matrices, same layout as pyn_elem_kle, equal to it to
 * round-off) or the fused integrate-and-assemble kernels (pyn_asm_desc.d_coef), which are bit-identical to each other. */
int pyn_elem_affine_coef(const pyn_elem* e, const double* d_coords, int64_t ne, double* d_coef, int64_t* h_nonaffine,
                         void* stream);
int pyn_elem_kle_affine(const pyn_elem* e, const double* d_coef, int64_t ne, double* d_K, double* d_Rw, double* d_Rd,
                        void* stream);
/* d_SrT [ne][3nn][2nn], d_DivSrT [ne][2nn][3nn], d_Curl [ne][nn][2nn], d_W [ne][nn] */
int pyn_elem_ops(const pyn_elem* e, const double* d_coords, int64_t ne,
                 double* d_SrT, double* d_DivSrT, double* d_Curl, double* d_W, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Mesh numbering on device.  Replaces the per-cell Python loops of DMPlexDom.getGlobalNodesFromCell /
 * IndicesManager.mapEntitiesToNodes (src/domain/dmplex.py:197-200, src/domain/indices.py:66-88),
 * getCellCornersCoords (dmplex.py:97-104) and computeFullCoordinates (dmplex.py:66-95) for the serial
 * DMPlexCreateBoxMesh numbering (dmplex.py:16).  Cells [cell0, cell0+ncells).
 * ---------------------------------------------------------------------------------------------------- */
int pyn_box_cell_nodes(int nx, int ny, int ngl, int64_t cell0, int64_t ncells,
                       int32_t* d_elem2node /* [ncells][ngl*ngl] */, void* stream);
int pyn_box_cell_corners(int nx, int ny, double x0, double y0, double x1, double y1, int64_t cell0, int64_t ncells,
                         double* d_coords /* [ncells][4][2] */, void* stream);
/* node coordinates [nnodes][2]: corner-bilinear map evaluated at the nodal GLL points (xi_nodal, ngl values) */
int pyn_box_node_coords(int nx, int ny, int ngl, double x0, double y0, double x1, double y1,
                        const double* h_xi_nodal, const int32_t* h_perm, double* d_xy, void* stream);
/* same for a strip of cell rows [fy0, fy0+ny_local) of the global nx x ny_global mesh, in the strip's own (local)
 * box numbering; coordinates are computed with the global mesh spacing so they are bit-identical to the serial ones */
int pyn_box_node_coords_strip(int nx, int ny_global, int64_t fy0, int ny_local, int ngl, double x0, double y0,
                              double x1, double y1, const double* h_xi_nodal, const int32_t* h_perm, double* d_xy,
                              void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Topology handle: node->element incidence and node adjacency (nodes sharing a cell), built once per mesh.
 * Replaces DMPlexDom.getMatIndices (src/domain/dmplex.py:305-333) and PETSc's MatAssembly pattern work.
 * elem ids in d_elem2node are LOCAL [0,ne); `elem_order` (NULL = identity) gives the global cell number used
 * for the summation order (ascending), so a partition sums in the same order as the serial cell loop.
 * ---------------------------------------------------------------------------------------------------- */
typedef struct pyn_topo pyn_topo;
int pyn_topo_create(const int32_t* d_elem2node, int64_t ne, int nn, int64_t nnodes, void* stream, pyn_topo** out);
int pyn_topo_destroy(pyn_topo* t);
/* nnz of the node adjacency, max elements per node, max adjacency row length */
int pyn_topo_info(const pyn_topo* t, int64_t* adj_nnz, int* max_valence, int* max_row);
/* device pointers to the node adjacency CSR (owned by the handle) */
int pyn_topo_adjacency(const pyn_topo* t, const int64_t** d_rowptr, const int32_t** d_cols);

/* ------------------------------------------------------------------------------------------------------
 * Deterministic global assembly into CSR (sorted columns, explicit zeros kept, duplicates added in ascending
 * cell order).  One descriptor describes one matrix of FreeSlip.buildKLEMats / NoSlipFreeSlip.buildKLEMats /
 * BaseProblem.buildOperators (src/cases/base_problem.py:499-552, :329-454, :132-140) together with
 * Mat.setIndices2One (src/matrices/mat_generator.py:113-118) and Operators.assembleAll (:172-190):
 *
 *   entry (row dof i of class ri, col dof j of class cj) is stored iff  pair_mask bit (ri*4+cj)  is set, or
 *   i == j and diag_classes bit ri is set;  its value is  rowscale[node(i)] * sign * sum_e E_e[li][lj]  (+
 *   diag_add[ri] on such a diagonal).  Classes: 0 free, 1 tangential-on-no-slip-wall, 2 prescribed.
 * ---------------------------------------------------------------------------------------------------- */
/* Boundary map (optional accelerator): the list of nodes that carry a non-zero row class or have a neighbour with
 * a non-zero column class.  All other nodes ("interior") keep either every entry or none, need no class lookups
 * and go through a streamlined kernel; boundary-only matrices (Krhs, Kfs, ...) touch only the listed nodes.
 * The class arrays are borrowed and must be the ones later passed in the descriptor. */
typedef struct pyn_bcmap pyn_bcmap;
int pyn_bcmap_create(const pyn_topo* t, const int8_t* d_rclass, int rdof, const int8_t* d_cclass, int cdof,
                     void* stream, pyn_bcmap** out);
int pyn_bcmap_destroy(pyn_bcmap* m);
int pyn_bcmap_info(const pyn_bcmap* m, int64_t* n_boundary_nodes);

typedef struct {
  int32_t rdof, cdof;        /* dofs per node on rows / columns: K 2x2, Rw 2x1, SrT 3x2, DivSrT 2x3, Curl 1x2 */
  const int8_t* d_rclass;    /* [nnodes*rdof] or NULL (all 0) */
  const int8_t* d_cclass;    /* [nnodes*cdof] or NULL (all 0) */
  uint32_t pair_mask;
  uint32_t diag_classes;
  double diag_add[4];
  double sign;
  const double* d_E;         /* [ne or 1][rdof*nn][cdof*nn] element matrices */
  int64_t E_stride;          /* elements between consecutive element matrices in doubles; 0 = one for all cells */
  const double* d_rowscale;  /* [nnodes] or NULL */
  const pyn_bcmap* bcmap;    /* NULL, or the map built from (d_rclass, rdof, d_cclass, cdof) */
  /* fused integrate-and-assemble on affine elements: when d_coef is set (pyn_elem_affine_coef) the entries of K_e
   * (which 0, 2x2 blocks), Rw_e (1) or Rd_e (2; 2x1 blocks) are evaluated while the rows are written and d_E is ignored */
  const double* d_coef;      /* [ne][18] or NULL */
  const pyn_elem* elem;      /* element handle the coefficients were computed with */
  int32_t which;
} pyn_asm_desc;

/* pattern: row pointers (exclusive scan done inside), total nnz returned on the host (synchronises) */
int pyn_asm_count(const pyn_topo* t, const pyn_asm_desc* desc, int64_t* d_indptr /* [nnodes*rdof+1] */,
                  int64_t* h_nnz, void* stream);
/* d_indices may be NULL (values only: re-assembly on a known pattern) ; d_data may be NULL (pattern only) */
int pyn_asm_fill(const pyn_topo* t, const pyn_asm_desc* desc, const int64_t* d_indptr,
                 int32_t* d_indices, double* d_data, void* stream);
/* Row templates (uniform meshes): when one element matrix serves every cell (E_stride == 0) and the topology found
 * few distinct row signatures, a values-only pyn_asm_fill computes each distinct interior row once and streams it to all
 * the nodes that share it (bit-identical to the gather kernel).  pyn_asm_set_templates(0) forces the gather kernel. */
int pyn_asm_set_templates(int enabled);
/* A values-only pyn_asm_fill runs its boundary-list kernel on a side stream of the topology handle, overlapped with the
 * interior rows, and joins it before returning.  pyn_asm_defer_join(1) postpones that join so that a batch of fills
 * (the eight matrices of base_problem.py:329-454) overlaps as a whole; the caller then MUST call pyn_asm_join(t, stream)
 * before anything reads the matrices, and pyn_asm_defer_join(0) afterwards. */
int pyn_asm_defer_join(int enabled);
int pyn_asm_join(const pyn_topo* t, void* stream);
int pyn_topo_signatures(const pyn_topo* t, int* nsig, int64_t* template_slots);
/* lumped weights  W[node] = sum_e W_e[local]  in ascending cell order  (mat_generator.py:168-170) */
int pyn_asm_weights(const pyn_topo* t, const double* d_We, int64_t We_stride, double* d_W, void* stream);

/* Generic deterministic COO -> CSR (Mat.setValues(ADD_VALUES) + assemble, mat_generator.py:14-17,157-166):
 * keys sorted by (row, col), duplicates added in input order.  Two calls: count then fill. */
int pyn_coo_count(int64_t n, const int32_t* d_rows, const int32_t* d_cols, int64_t nrows, int64_t ncols,
                  int64_t* d_indptr, int64_t* h_nnz, void** scratch /* opaque, freed by fill */, void* stream);
int pyn_coo_fill(void* scratch, const double* d_vals, const int64_t* d_indptr, int32_t* d_indices, double* d_data,
                 void* stream);

/* C = A + alpha*B on the union pattern (petsc4py `K + Kfs`, base_problem.py:318) */
int pyn_csr_add_count(int64_t nrows, const int64_t* d_aptr, const int32_t* d_aind,
                      const int64_t* d_bptr, const int32_t* d_bind, int64_t* d_cptr, int64_t* h_nnz, void* stream);
int pyn_csr_add_fill(int64_t nrows, const int64_t* d_aptr, const int32_t* d_aind, const double* d_aval, double alpha,
                     const int64_t* d_bptr, const int32_t* d_bind, const double* d_bval,
                     const int64_t* d_cptr, int32_t* d_cind, double* d_cval, void* stream);
int pyn_csr_diagonal(int64_t nrows, const int64_t* d_ptr, const int32_t* d_ind, const double* d_val,
                     double* d_diag, void* stream);
/* The same for a pattern that is re-used (values-only re-assembly): pyn_csr_diag_offsets finds, once, the offset of the
 * diagonal entry inside every row (-1: the row stores none); pyn_csr_diagonal_at then gathers diag[r] =
 * val[ptr[r] + off[r]] without searching. */
int pyn_csr_diag_offsets(int64_t nrows, const int64_t* d_ptr, const int32_t* d_ind, int32_t* d_off, void* stream);
int pyn_csr_diagonal_at(int64_t nrows, const int64_t* d_ptr, const int32_t* d_off, const double* d_val, double* d_diag,
                        void* stream);
/* drop stored zeros (compressed operator pattern for the timed SpMV): count then fill */
int pyn_csr_compress_count(int64_t nrows, const int64_t* d_ptr, const double* d_val, int64_t* d_optr, int64_t* h_nnz,
                           void* stream);
int pyn_csr_compress_fill(int64_t nrows, const int64_t* d_ptr, const int32_t* d_ind, const double* d_val,
                          const int64_t* d_optr, int32_t* d_oind, double* d_oval, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * SpMV  y = A x  (PETSc Mat.mult / `Mat * Vec`, base_problem.py:219,229,232,323-327,481) with fused epilogues.
 * ---------------------------------------------------------------------------------------------------- */
typedef enum {
  PYN_SPMV_SET = 0,        /* y = A x */
  PYN_SPMV_ADD = 1,        /* y += A x */
  PYN_SPMV_SCALE = 2,      /* y = p0 * (A x)                                 (base_problem.py:229-230) */
  PYN_SPMV_SRT_VTENSV = 3  /* y = p0*(A x) - p1*VtensV(aux)  rows = 3*nodes   (base_problem.py:217-223) */
} pyn_spmv_mode;
int pyn_spmv(int64_t nrows, const int64_t* d_ptr, const int32_t* d_ind, const double* d_val, int64_t nnz,
             const double* d_x, double* d_y, int mode, double p0, double p1, const double* d_aux, void* stream);

/* Planned SpMV ("CSR-stream"): the plan cuts the matrix into chunks of consecutive non-zeros snapped to row
 * boundaries, once per sparsity pattern; the kernel then streams values/indices fully coalesced whatever the row
 * lengths.  Same result contract and modes as pyn_spmv.  `d_ptr` is borrowed and must outlive the plan. */
typedef struct pyn_spmv_plan pyn_spmv_plan;
int pyn_spmv_plan_create(int64_t nrows, const int64_t* d_ptr, int64_t nnz, void* stream, pyn_spmv_plan** out);
int pyn_spmv_plan_destroy(pyn_spmv_plan* plan);
int pyn_spmv_planned(const pyn_spmv_plan* plan, const int32_t* d_ind, const double* d_val, const double* d_x,
                     double* d_y, int mode, double p0, double p1, const double* d_aux, void* stream);

/* Element ("template") operator for uniform meshes: the same y = A x without streaming the matrix values -- the
 * matrix-free form of MatMult on the matrices of base_problem.py:329-454,499-552 / mat_generator.py:157-190.  Built
 * from the descriptor a matrix was assembled with (ONE element matrix for all cells, boundary map) plus the assembled
 * CSR, whose rows serve the boundary-affected nodes only.  Interior nodes: phase 1 multiplies the element matrix
 * (resident in shared memory) with the gathered dofs of every element, phase 2 sums the element results per node in
 * ascending cell order (deterministic, no atomics).  The CSR may be any matrix that equals the descriptor's matrix on
 * the interior rows (e.g. K + Kfs for K).  The descriptor's element matrix is copied, everything else is borrowed.
 * Same modes as pyn_spmv.  pyn_tmat_supported: 1 when an element matrix of rdof*nn x cdof*nn fits the kernel's shared
 * memory budget (ngl <= 7 for K), else the caller keeps the CSR stream. */
typedef struct pyn_tmat pyn_tmat;
int pyn_tmat_create(const pyn_topo* t, const pyn_asm_desc* desc, const int64_t* d_indptr, const int32_t* d_indices,
                    const double* d_data, void* stream, pyn_tmat** out);
int pyn_tmat_destroy(pyn_tmat* m);
int pyn_tmat_mult(const pyn_tmat* m, const double* d_x, double* d_y, int mode, double p0, double p1, const double* d_aux,
                  void* stream);
int64_t pyn_tmat_bytes(const pyn_tmat* m);
int pyn_tmat_supported(int nn, int rdof, int cdof);

/* ------------------------------------------------------------------------------------------------------
 * Immersed-boundary coupling (ImmersedBoundaryStatic / Dynamic, src/cases/immersed_boundary.py)
 * ------------------------------------------------------------------------------------------------------
 * pyn_ibm_collect replaces collectNodes + computeClose + computeDirac (immersed_boundary.py:258-277, 318-338 and, for
 * mode 1, :391-402): for every Lagrangian node (d_lagxy [nlag][2]) the candidate Euler nodes (d_cand, ascending ids into
 * d_nodexy [nnodes][2]) that are "close" -- mode 0: |dx / dl| < 3 and |dy / dl| < 3; mode 1: |dx / dl| < 2 and
 * |dy| < 1 + 2 dl -- with the discrete delta value prod_r phi(|r| / h) / h (immersed_body.py:259-264; delta_kind 0
 * fourGrid, 1 threeGrid, 2 linear, :398-421).  Two passes over (marker, chunk) pairs, marker-major, with nchunks =
 * pyn_ibm_chunks(ncand, NULL) chunks of the candidate list per marker: d_ptr == NULL writes the counts [nlag * nchunks]
 * into d_count_or_nodes; with their prefix sums d_ptr [nlag * nchunks + 1] it writes the node ids and delta values
 * (ascending node id; the list of marker l starts at d_ptr[l * nchunks]).
 * pyn_ibm_gram replaces H.matMult(S) of buildIBMMatrix (:246): the Lagrangian system matrix for one velocity
 * component, entry (i, j) = sum over the common Euler nodes e, ascending, of (d_i(e) a)(d_j(e) b) with a = h^2 and
 * b = dl h the scale factors of H and S (:243-245); the same two-pass protocol over (marker, block of 32 partner
 * markers) pairs: counts [nlag * ceil(nlag / 32)], then the partner indices j and the values at their prefix sums. */
int pyn_ibm_chunks(int64_t n, int64_t* chunk);
int pyn_ibm_collect(int64_t nlag, const double* d_lagxy, int64_t ncand, const int32_t* d_cand, const double* d_nodexy,
                    double dl, double h, int mode, int delta_kind, const int64_t* d_ptr, int32_t* d_count_or_nodes,
                    double* d_dirac, void* stream);
int pyn_ibm_gram(int64_t nlag, const int64_t* d_ptr, const int32_t* d_nodes, const double* d_dirac, double a, double b,
                 const int64_t* d_aptr, int32_t* d_count_or_cols, double* d_avals, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Vectors (PETSc Vec calls of the path: base_problem.py:222-230,234-252; cavity.py:53-81; custom_func.py:76-83)
 * ---------------------------------------------------------------------------------------------------- */
int pyn_vec_set(int64_t n, double a, double* d_x, void* stream);
int pyn_vec_scale(int64_t n, double a, double* d_x, void* stream);
int pyn_vec_axpy(int64_t n, double a, const double* d_x, double* d_y, void* stream);          /* y += a x */
int pyn_vec_aypx(int64_t n, double a, const double* d_x, double* d_y, void* stream);          /* y = x + a y */
int pyn_vec_waxpy(int64_t n, double a, const double* d_x, const double* d_y, double* d_w, void* stream); /* w = a x + y */
int pyn_vec_maxpy(int64_t n, int m, const double* h_coef, const double* const* h_xs, const double* d_y0,
                  double* d_y, void* stream);                          /* y = y0 + sum c_j x_j ; y0 NULL = 0 */
int pyn_vec_dot(int64_t n, const double* d_x, const double* d_y, double* h_out, void* stream); /* synchronises */
int pyn_vec_norm(int64_t n, const double* d_x, int kind /*2: l2, 3: inf*/, double* h_out, void* stream);
/* pyn_vec_dot for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  Bit-identical to pyn_vec_dot, but the
 * per-CTA partial sums live in a file-scope device array of PYN_VEC_DOT_NOALLOC_MAX_PARTS doubles (16 KiB, held by every
 * GPU that loads the library) and the result is written to the caller's DEVICE double d_out; the call allocates and
 * frees nothing and does not synchronise.  n > PYN_VEC_DOT_NOALLOC_MAX_N is rejected (PYN_ERR_ARG).  All calls share the
 * array: calls that overlap on the GPU (different streams) race. */
#define PYN_VEC_DOT_NOALLOC_MAX_PARTS 2048
#define PYN_VEC_DOT_NOALLOC_MAX_N ((int64_t)1 << 28)
int pyn_vec_dot_noalloc(int64_t n, const double* d_x, const double* d_y, double* d_out, void* stream);
int pyn_vec_scatter_set(int64_t nidx, const int64_t* d_idx, const double* d_vals, double* d_x, void* stream);
int pyn_vec_scatter_const(int64_t nidx, const int64_t* d_idx, double val, double* d_x, void* stream);
int pyn_vtensv(int64_t nnodes, const double* d_vel, double* d_out, void* stream);  /* base_problem.py:234-252 */
/* Taylor-Green analytic fields on a node list (custom_func.py:76-83,173-193); idx NULL = all nodes */
int pyn_taylor_green(int64_t nidx, const int64_t* d_idx, const double* d_xy, double nu, double t,
                     double* d_vel /* nullable */, double* d_vort /* nullable */, void* stream);
/* TSAdapt-basic weighted RMS error norm: sqrt(mean((e_i/(atol+rtol*max(|y_i|,|ynew_i|)))^2)), synchronises */
int pyn_vec_wrms(int64_t n, const double* d_err, const double* d_y, const double* d_ynew, double atol, double rtol,
                 double* h_out, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Krylov solvers with PETSc semantics (KspSolver, src/solver/ksp_solver.py:9-19; use sites
 * base_problem.py:276,318,323-327,481): zero initial guess, left preconditioning, preconditioned residual
 * norm, KSPConvergedDefault.  reason: 2 rtol, 3 atol, -3 max_it, -4 dtol, -8 indefinite.
 * nonzero_guess (PETSc's -ksp_initial_guess_nonzero): d_x holds x0 on entry, r0 = b - A x0, and rtol is
 * measured against |M^-1 b| (the initial residual when b = 0), as KSPConvergedDefault does.
 * ---------------------------------------------------------------------------------------------------- */
typedef struct pyn_ksp pyn_ksp;
typedef struct {
  int32_t type;        /* 0 cg, 1 gmres */
  int32_t pc;          /* 0 none, 1 jacobi, 2 multigrid V-cycle (cg only; attach with pyn_ksp_set_mg) */
  double rtol, atol, dtol;
  int32_t max_it;
  int32_t restart;     /* gmres */
  int32_t check_every; /* iterations between host polls of the device convergence flag (0 = default) */
  int32_t nonzero_guess; /* start from the contents of d_x */
} pyn_ksp_opts;
typedef struct {
  int32_t its, reason;
  double rnorm0, rnorm;
} pyn_ksp_result;
int pyn_ksp_create(int64_t nrows, const int64_t* d_ptr, const int32_t* d_ind, const double* d_val, int64_t nnz,
                   const pyn_ksp_opts* opts, void* stream, pyn_ksp** out);
int pyn_ksp_solve(pyn_ksp* k, const double* d_b, double* d_x, pyn_ksp_result* res, void* stream); /* synchronises */
int pyn_ksp_destroy(pyn_ksp* k);
/* device work done by the last solve: SpMV launches and total kernel launches (for the roofline accounting) */
int pyn_ksp_stats(const pyn_ksp* k, int64_t* spmv_launches, int64_t* kernel_launches);
/* CG and GMRES apply the operator through the template operator instead of the CSR stream (NULL: back to the CSR) */
int pyn_ksp_set_tmat(pyn_ksp* k, const pyn_tmat* m);

/* ------------------------------------------------------------------------------------------------------
 * Geometric p/h-multigrid V-cycle as the preconditioner of CG (`-pc_type mg`; SURVEY 8f-1: the reference's answer to
 * the cost of the solve is a factor-once LU, src/solver/ksp_solver.py:13-16).  The caller (pynama_b200/solver/
 * multigrid.py) builds the levels -- the rediscretised K of the same mesh at lower orders, then of bilinear elements on
 * meshes coarsened by two -- with the library's own integration / assembly kernels and hands over, per level: the
 * operator (element operator and/or planned CSR), 1/diag on the unknowns (0 on prescribed dofs), an upper bound of
 * the spectrum of D^-1 A; per level pair: interpolation P and restriction R = P^T as planned CSR; and the dense
 * inverse of the coarsest operator.  Smoother: Chebyshev of `degree` on [lmax / ratio, lmax], before and after the
 * coarse correction (symmetric cycle).  Everything is borrowed; work vectors are allocated on first use.
 * ---------------------------------------------------------------------------------------------------- */
typedef struct pyn_mg pyn_mg;
int pyn_mg_create(int nlevels, int degree, double ratio, pyn_mg** out);
int pyn_mg_set_level(pyn_mg* g, int level, int64_t n, const pyn_tmat* tmat, const pyn_spmv_plan* plan,
                     const int32_t* d_ind, const double* d_val, const double* d_dinv, double lmax);
int pyn_mg_set_transfer(pyn_mg* g, int level, const pyn_spmv_plan* Pplan, const int32_t* d_Pind, const double* d_Pval,
                        const pyn_spmv_plan* Rplan, const int32_t* d_Rind, const double* d_Rval);
int pyn_mg_set_coarse_inverse(pyn_mg* g, const double* d_Ainv);   /* [n_c][n_c], row major; NULL: smoother only */
int pyn_mg_apply(pyn_mg* g, const double* d_r, double* d_z, void* stream);   /* z = V(r); 0 on prescribed dofs */
int pyn_mg_destroy(pyn_mg* g);
int pyn_ksp_set_mg(pyn_ksp* k, pyn_mg* g);
/* Partitioned mesh: level 0 is this rank's strip (halo exchange before every fine product, restricted residual summed
 * over the ranks); all coarser levels are replicated on every rank.  Declared after the multi-GPU section below. */

/* ------------------------------------------------------------------------------------------------------
 * Multi-GPU: one process per GPU, the mesh cut into strips of cell rows (the reference: DMPlex.distribute(),
 * src/domain/dmplex.py:29; MatStash / VecScatter / MPI_Allreduce inside PETSc).  NCCL is bound at run time.
 *   - pyn_comm: NCCL communicator.  Rank 0 makes the 128-byte id, the host side broadcasts it (torch.distributed).
 *   - pyn_halo: per neighbour, the local node ids to send (owned nodes the neighbour holds as ghosts) and to receive
 *     into (ghost nodes), both ordered by global node id; one grouped ncclSend/ncclRecv per exchange.
 *   - pyn_ksp_set_dist: CG then refreshes the ghosts of the search direction before every SpMV and all-reduces its
 *     three scalars; `d_owned_rows` (1 = row owned by this rank) masks the right-hand side.
 * ---------------------------------------------------------------------------------------------------- */
typedef struct pyn_comm pyn_comm;
typedef struct pyn_halo pyn_halo;
int pyn_comm_unique_id(char* id128);
int pyn_comm_create(const char* id128, int rank, int world, pyn_comm** out);
int pyn_comm_destroy(pyn_comm* c);
int pyn_comm_info(const pyn_comm* c, int* rank, int* world, int* nccl_version);
int pyn_comm_allreduce_sum(const pyn_comm* c, double* d_buf, int n, void* stream);
int pyn_halo_create(const pyn_comm* c, int nnbr, const int32_t* h_nbr_rank, const int64_t* h_send_cnt,
                    const int64_t* const* d_send_idx, const int64_t* h_recv_cnt, const int64_t* const* d_recv_idx,
                    pyn_halo** out);
int pyn_halo_destroy(pyn_halo* h);
/* Peer memory: one region per rank, mapped by every other rank of the box through CUDA IPC (NVLink P2P).  Once attached
 * (pyn_comm_set_peer) the all-reduces of <= 7 scalars become a one-warp kernel that stores the rank's partial into
 * every peer's slot and sums the slots in rank order, and the halo exchange packs straight into the neighbour's
 * mailbox and hands over with an epoch flag -- no NCCL launch on the Krylov iteration's critical path.  The caller
 * gathers the 64-byte handles of all ranks (rank-major) with its rendezvous layer. */
typedef struct pyn_peer pyn_peer;
int pyn_peer_create(int rank, int world, int64_t mailbox_doubles, pyn_peer** out);
int pyn_peer_handle(const pyn_peer* p, char* handle64);
int pyn_peer_open(pyn_peer* p, const char* handles /* [world][64] */);
int pyn_peer_destroy(pyn_peer* p);
int pyn_comm_set_peer(pyn_comm* c, pyn_peer* p);
/* d_vec: local vector with ncomp (1..3) dofs per node; ghost entries are overwritten with their owners' values */
int pyn_halo_exchange(const pyn_halo* h, double* d_vec, int ncomp, void* stream);
int64_t pyn_halo_bytes(const pyn_halo* h, int ncomp);
int pyn_ksp_set_dist(pyn_ksp* k, const pyn_comm* comm, const pyn_halo* halo, int ncomp, const uint8_t* d_owned_rows);
int pyn_mg_set_dist(pyn_mg* g, const pyn_comm* comm, const pyn_halo* halo, int ncomp);   /* level 0 */
/* levels 0 .. k (the fine mesh at decreasing order) may each be partitioned with their own halo plan; the rest is replicated */
int pyn_mg_set_level_dist(pyn_mg* g, int level, const pyn_comm* comm, const pyn_halo* halo, int ncomp);

#ifdef __cplusplus
}
#endif
#endif /* PYNAMA_B200_H */
