// Topology handle shared between topo.cu and assemble.cu.
#pragma once
#include "common.cuh"

struct pyn_topo {
  int64_t ne, nnodes;
  int nn;
  const int* e2n;      // [ne][nn] borrowed from the caller (must outlive the handle)
  int64_t* n2e_ptr;    // [nnodes+1]
  int* n2e_elem;       // element ids ascending per node
  short* n2e_loc;      // local index of the node inside that element
  int64_t* adj_ptr;    // [nnodes+1]
  int* adj_cols;       // sorted neighbour nodes (including the node itself)
  int64_t* inv_ptr;    // [nnodes+1] offsets into `inv` (sum of valence * row length)
  uint8_t* inv;        // inverse position map: inv[inv_ptr[r] + k*L + p] = local index, inside the k-th element of
                       // node r, of the p-th adjacency column of r (0xFF: that column is not a node of the element)
  int64_t adj_nnz, inv_len;
  int max_valence, max_row;
  // Row signatures (uniform-mesh fast path of the assembly).  Two nodes whose incidence lists (local indices inside
  // their elements) and inverse position maps are byte-identical get identical rows from ONE element matrix, so the
  // rows are computed once per signature ("template") and streamed to every node that carries it.
  int* sig_id;        // [nnodes] signature of each node
  int nsig;           // 0 = signatures not available (irregular mesh or hash collision)
  int* sig_rep;       // [nsig] representative (lowest-numbered) node
  int64_t* sig_ptr;   // [nsig+1] offsets of the templates in adjacency slots
  int64_t sig_slots;
  double* sig_tmpl;    // [sig_slots * 9] scratch for the row templates of the matrix being assembled (lazy)
  // Side stream of the values-only assembly: the boundary-list kernels (short, dependent-latency bound) run there
  // while the caller's stream writes the interior rows.  `pending`: side work not yet joined (pyn_asm_join).
  cudaStream_t side;
  cudaEvent_t ev_fork, ev_join;
  int pending;
  // Inside a batch of fills (pyn_asm_defer_join) the row-template kernels of the later matrices run on a second side
  // stream, forked once per batch after the element matrices exist, so that they overlap the store stream of the
  // previous matrix instead of sitting between two stream kernels; a ring of template buffers keeps them apart.
  cudaStream_t side2;
  cudaEvent_t ev_batch, ev_tmpl[4];
  double* sig_ring[4];
  int batch_forked, ring_next;
};

struct pyn_bcmap {
  int64_t nnodes, nlist;
  uint8_t* flags;  // [nnodes] 1 = boundary-affected node
  int* list;       // the flagged nodes, ascending
  cudaStream_t st; // stream the buffers were allocated on (stream-ordered allocator) and are released on
  // Signature tiles (lazy, structured meshes only): the un-flagged nodes grouped by row signature (ascending node id
  // inside a group) and cut into tiles of <= 32 nodes of ONE signature: tile i = nodes sig_nodes[tile_start[i] ..
  // + tile_cnt[i]) of signature tile_sig[i].  Used by the signature-grouped fused assembly kernel (assemble.cu).
  int* sig_nodes;
  int *tile_sig, *tile_start, *tile_cnt;
  // What the kernel reads (built from the above, which are then released): per tile 4 ints {cnt | deg << 8 | L << 16,
  // local row index of the representative in its elements (4 bytes), offset of its inverse map (2 ints)}; per tile and
  // lane a slot {node, its elements} of 2 / 4 / 8 ints for the valence classes 1 / 2 / <= 4 (lanes beyond cnt repeat
  // lane 0); class c starts at int tile_slot_base[c] of the slot table
  int* tile_desc;
  int* tile_slot;
  int64_t ntiles;
  // tiles are sorted by class, then by first node.  Class = 2 * valence index (1, 2, <= 4 incident elements) + 1 when the
  // tile takes the line-aligned kernel (rows long enough and not whole sectors; never for valence 4): PYN_SIG_CLASSES = 5
  int64_t tile_slot_base[5];
  int64_t tile_class_end[5];
  int tiles_state;   // 0 not built, 1 built, -1 unavailable
  const int64_t* tiles_indptr;   // the matrix whose row starts the tiles were grouped by
};

namespace pyn {
struct TopoView {
  int64_t ne, nnodes;
  int nn;
  const int* e2n;
  const int64_t* n2e_ptr;
  const int* n2e_elem;
  const short* n2e_loc;
  const int64_t* adj_ptr;
  const int* adj_cols;
  const int64_t* inv_ptr;
  const uint8_t* inv;
  const int* sig_id;
  const int* sig_rep;
  const int64_t* sig_ptr;
  int nsig;
};
static inline TopoView view_of(const pyn_topo* t) {
  TopoView v;
  v.ne = t->ne; v.nnodes = t->nnodes; v.nn = t->nn; v.e2n = t->e2n;
  v.n2e_ptr = t->n2e_ptr; v.n2e_elem = t->n2e_elem; v.n2e_loc = t->n2e_loc;
  v.adj_ptr = t->adj_ptr; v.adj_cols = t->adj_cols; v.inv_ptr = t->inv_ptr; v.inv = t->inv;
  v.sig_id = t->sig_id; v.sig_rep = t->sig_rep; v.sig_ptr = t->sig_ptr; v.nsig = t->nsig;
  return v;
}
int scan_counts(const int* d_cnt, int64_t n, int64_t* d_ptr, int64_t* h_total, cudaStream_t st);
}  // namespace pyn
