#include <stdlib.h>
// Deterministic global assembly into CSR.
//
// Row-gather formulation (no atomics): one warp owns one mesh node = `rdof` consecutive matrix rows.  It walks the
// node's elements in ASCENDING cell order and adds each element block into shared-memory accumulators indexed by
// the node's sorted adjacency list, then writes the rows out compacted by the boundary-condition masks.  The
// summation order is therefore exactly the reference's serial cell loop (base_problem.py:337,504,135) and the
// result is bitwise reproducible run to run.
#include <algorithm>
#include <vector>

#include <cub/cub.cuh>

#include "common.cuh"
#include "topo.cuh"
#include "tmat.cuh"
#include "elem.cuh"

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
  const uint8_t* flags;  // bcmap: 1 = boundary-affected node (NULL: no map)
  // fused integrate-and-assemble on affine elements (elem_affine.cuh): entries are evaluated from the per-element
  // coefficients instead of being read from materialised element matrices (coef == NULL: E is used)
  const double* coef;    // [ne][PYN_AFF_NCOEF]
  pyn::AffTables aff;
  int which;             // 0 K (2x2 blocks), 1 Rw, 2 Rd (2x1 blocks)
  // signature tiles of the boundary map (structured meshes; NULL: not available)
  const int *tile_desc, *tile_slot;
  int64_t ntiles;
  int64_t tile_class_end[5], tile_slot_base[5];
};

// element block (rows of local node lr, columns of local node l) of one element from the affine closed form:
// cf = the element's coefficients, rec = the product record of the pair (lr, l)
template <int RD, int CD>
__device__ __forceinline__ void affine_entry(const DescView& d, const double* __restrict__ cf,
                                             const double* __restrict__ rec, double (&v)[RD][CD]) {
  if constexpr (RD == 2 && CD == 2) {
    pyn::affine_block_K(cf, rec, v);
  } else if constexpr (RD == 2 && CD == 1) {
    pyn::affine_block_R(cf + 13, rec, d.which, d.aff.alpha_w, d.aff.alpha_d, v);
  }
}

// One warp per mesh node (= RD consecutive matrix rows); lanes stride over the node's sorted adjacency row.  Every
// lane owns whole output blocks: for adjacency position p it gathers, in ascending cell order, the RD x CD block of
// each incident element that contains column node p (inverse position map of the topology handle), so there are no
// shared-memory accumulators, no atomics and no intra-warp ordering hazards.  COUNT only counts kept entries.
template <int RD, int CD, bool COUNT, bool AFF = false>
__global__ void __launch_bounds__(256)
asm_gather_kernel(pyn::TopoView t, DescView d, const int* __restrict__ list, int64_t nlist, int* __restrict__ rowcnt,
                  const int64_t* __restrict__ indptr, int* __restrict__ indices, double* __restrict__ data) {
  const int lane = threadIdx.x & 31;
  const int64_t warp0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  const bool values = (!COUNT) && (data != nullptr);
  const unsigned lt = (1u << lane) - 1u;
  const int ldE = CD * t.nn;
  const int64_t nwork = list ? nlist : t.nnodes;
  for (int64_t wi = warp0; wi < nwork; wi += nwarps) {
    const int64_t r = list ? (int64_t)list[wi] : wi;
    const int64_t a0 = t.adj_ptr[r], b0 = t.n2e_ptr[r];
    const int L = (int)(t.adj_ptr[r + 1] - a0), deg = (int)(t.n2e_ptr[r + 1] - b0);
    const int64_t iv0 = values ? t.inv_ptr[r] : 0;
    int my_e = 0, my_lr = 0;
    if (values && lane < deg) {
      my_e = t.n2e_elem[b0 + lane];
      my_lr = t.n2e_loc[b0 + lane];
    }
    int rca[RD];
    int64_t base[RD];
    int running[RD];
#pragma unroll
    for (int a = 0; a < RD; ++a) {
      rca[a] = d.rclass ? d.rclass[r * RD + a] : 0;
      base[a] = COUNT ? 0 : indptr[r * RD + a];
      running[a] = 0;
    }
    const double rs = (values && d.rowscale) ? d.rowscale[r] : 1.0;
    for (int p0 = 0; p0 < L; p0 += 32) {
      const int p = p0 + lane;
      const bool valid = p < L;
      const int c = valid ? t.adj_cols[a0 + p] : 0;
      int cc[CD];
#pragma unroll
      for (int b = 0; b < CD; ++b) cc[b] = (valid && d.cclass) ? d.cclass[(int64_t)c * CD + b] : 0;
      double v[RD][CD];
#pragma unroll
      for (int a = 0; a < RD; ++a)
#pragma unroll
        for (int b = 0; b < CD; ++b) v[a][b] = 0.0;
      if (values) {
        for (int k = 0; k < deg; ++k) {  // ascending cell order: the reference's serial cell loop
          const int e = __shfl_sync(0xffffffffu, my_e, k), lr = __shfl_sync(0xffffffffu, my_lr, k);
          const int l = valid ? (int)t.inv[iv0 + (int64_t)k * L + p] : 0xFF;
          if (l != 0xFF) {
            if constexpr (AFF) {
              double blk[RD][CD];
              const double* rec = CD == 2 ? d.aff.pairK + (int64_t)(lr * t.nn + l) * pyn::PYN_AFF_PK
                                          : d.aff.pairR + (int64_t)(lr * t.nn + l) * pyn::PYN_AFF_PR;
              affine_entry<RD, CD>(d, d.coef + (int64_t)e * PYN_AFF_NCOEF, rec, blk);
#pragma unroll
              for (int a = 0; a < RD; ++a)
#pragma unroll
                for (int b = 0; b < CD; ++b) v[a][b] += d.sign * blk[a][b];
            } else {
            const double* Ee = d.E + (d.E_stride ? (int64_t)e * d.E_stride : 0) + (int64_t)lr * RD * ldE + l * CD;
#pragma unroll
            for (int a = 0; a < RD; ++a)
#pragma unroll
              for (int b = 0; b < CD; ++b) v[a][b] += d.sign * Ee[a * ldE + b];
            }
          }
        }
      }
#pragma unroll
      for (int a = 0; a < RD; ++a) {
        const bool dgrow = (RD == CD) && ((d.diag_classes >> rca[a]) & 1u);
        bool keep[CD], pa[CD];
        int before = 0, mine = 0, total = 0;
#pragma unroll
        for (int b = 0; b < CD; ++b) {
          pa[b] = (d.pair_mask >> (rca[a] * 4 + cc[b])) & 1u;
          const bool dg = dgrow && (c == r) && (a == b);
          keep[b] = valid && (pa[b] || dg);
          const unsigned m = __ballot_sync(0xffffffffu, keep[b]);
          before += __popc(m & lt);
          total += __popc(m);
        }
        if (!COUNT) {
#pragma unroll
          for (int b = 0; b < CD; ++b) {
            if (keep[b]) {
              const int64_t pos = base[a] + running[a] + before + mine;
              if (indices) indices[pos] = c * CD + b;
              if (values) {
                double x = pa[b] ? v[a][b] : 0.0;
                if (d.rowscale) x = rs * x;
                if (dgrow && (c == r) && (a == b))
                  x += (rca[a] == 0) ? d.diag_add[0] : (rca[a] == 1) ? d.diag_add[1] : (rca[a] == 2) ? d.diag_add[2] : d.diag_add[3];
                data[pos] = x;
              }
              ++mine;
            }
          }
        }
        running[a] += total;
      }
    }
    if (COUNT && lane == 0) {
#pragma unroll
      for (int a = 0; a < RD; ++a) rowcnt[r * RD + a] = running[a];
    }
  }
}

// Interior nodes (bcmap flag 0): every class is 0, so the node keeps either all its L*CD entries per row or none;
// no class lookups, no compaction, positions are p*CD+b.  Launched only when the (free,free) pair is kept.
//
// A CTA takes ASM_NB consecutive nodes: their headers (row pointers, incidence lists, output bases) are staged in
// shared memory with one coalesced round of loads, then the CTA's threads share the nodes' adjacency positions --
// which are contiguous in the adjacency CSR -- one output block per thread and step.  Every thread's work is then a
// short independent chain (inverse-map byte -> element-matrix entries from L1 -> coalesced store), so the kernel is
// limited by the store stream instead of by per-node latency.
constexpr int ASM_NB = 32;     // nodes per CTA batch
constexpr int ASM_MAXDEG = 8;  // nodes with more incident elements are routed to the generic kernel by the bcmap
constexpr int ASM_IVCAP = 12288;  // bytes of the inverse position map staged in shared memory per batch

template <int RD, int CD, bool VEC>
__global__ void __launch_bounds__(256)
asm_interior_kernel(pyn::TopoView t, DescView d, const int64_t* __restrict__ indptr, int* __restrict__ indices,
                    double* __restrict__ data) {
  __shared__ int64_t s_adj[ASM_NB + 1], s_inv[ASM_NB + 1], s_base[ASM_NB * RD], s_eoff[ASM_NB * ASM_MAXDEG];
  __shared__ __align__(16) uint8_t s_ivb[ASM_IVCAP + 32];
  __shared__ int s_n2e[ASM_NB + 1];
  __shared__ double s_rs[ASM_NB];
  __shared__ uint8_t s_flag[ASM_NB];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int ldE = CD * t.nn;
  const int64_t nbatch = (t.nnodes + ASM_NB - 1) / ASM_NB;
  for (int64_t batch = blockIdx.x; batch < nbatch; batch += gridDim.x) {
    const int64_t r0 = batch * ASM_NB;
    const int nb = (int)min((int64_t)ASM_NB, t.nnodes - r0);
    const int64_t e0 = t.n2e_ptr[r0];
    if (tid <= nb) {
      s_adj[tid] = t.adj_ptr[r0 + tid];
      s_n2e[tid] = (int)(t.n2e_ptr[r0 + tid] - e0);
      s_inv[tid] = t.inv_ptr[r0 + tid];
    }
    if (tid < nb) {
      s_flag[tid] = d.flags[r0 + tid];
      s_rs[tid] = d.rowscale ? d.rowscale[r0 + tid] : 1.0;
    }
    if (tid >= 64 && tid < 64 + nb * RD) s_base[tid - 64] = indptr[r0 * RD + (tid - 64)];
    __syncthreads();
    // the batch's slice of the inverse position map is contiguous: stage it with 16-byte loads (aligned superset)
    const uint8_t* ivsrc = t.inv + s_inv[0];
    const int mis = (int)(reinterpret_cast<uintptr_t>(ivsrc) & 15);
    const int64_t ivlen = s_inv[nb] - s_inv[0];
    const bool staged = ivlen + mis <= ASM_IVCAP;
    if (staged && data) {
      const uint4* src = reinterpret_cast<const uint4*>(ivsrc - mis);
      const int n16 = (int)((ivlen + mis + 15) >> 4);
      for (int i = tid; i < n16; i += 256) reinterpret_cast<uint4*>(s_ivb)[i] = src[i];
    }
    {  // element-matrix row offsets of every (node, incident element): 8 threads per node
      const int j = tid >> 3, k = tid & 7;
      if (j < nb && !s_flag[j] && k < s_n2e[j + 1] - s_n2e[j]) {
        const int64_t q = e0 + s_n2e[j] + k;
        s_eoff[j * ASM_MAXDEG + k] = (d.E_stride ? (int64_t)t.n2e_elem[q] * d.E_stride : 0) +
                                     (int64_t)t.n2e_loc[q] * RD * ldE;
      }
    }
    __syncthreads();
    for (int j = w; j < nb; j += 8) {  // one warp per node, lanes over its adjacency positions
      if (s_flag[j]) continue;
      const int L = (int)(s_adj[j + 1] - s_adj[j]), deg = s_n2e[j + 1] - s_n2e[j];
      const uint8_t* ivj = staged ? s_ivb + mis + (int)(s_inv[j] - s_inv[0]) : t.inv + s_inv[j];
      const double rs = s_rs[j];
      for (int p = lane; p < L; p += 32) {
        double v[RD][CD];
#pragma unroll
        for (int a = 0; a < RD; ++a)
#pragma unroll
          for (int b = 0; b < CD; ++b) v[a][b] = 0.0;
        if (data) {
          for (int k = 0; k < deg; ++k) {  // ascending cell order
            const int l = ivj[k * L + p];
            if (l != 0xFF) {
              const double* Ee = d.E + s_eoff[j * ASM_MAXDEG + k] + l * CD;
#pragma unroll
              for (int a = 0; a < RD; ++a) {
                if (VEC) {  // CD == 2, 16-byte aligned element matrices
                  const double2 e2 = *reinterpret_cast<const double2*>(Ee + a * ldE);
                  v[a][0] += d.sign * e2.x;
                  v[a][CD - 1] += d.sign * e2.y;
                } else {
#pragma unroll
                  for (int b = 0; b < CD; ++b) v[a][b] += d.sign * Ee[a * ldE + b];
                }
              }
            }
          }
#pragma unroll
          for (int a = 0; a < RD; ++a) {
            const int64_t base = s_base[j * RD + a];
            double* o = data + base + p * CD;
#pragma unroll
            for (int b = 0; b < CD; ++b)
              if (d.rowscale) v[a][b] = rs * v[a][b];
            if (CD == 2 && ((base & 1) == 0)) {
              *reinterpret_cast<double2*>(o) = make_double2(v[a][0], v[a][CD - 1]);
            } else {
#pragma unroll
              for (int b = 0; b < CD; ++b) o[b] = v[a][b];
            }
          }
        }
        if (indices) {
          const int c = t.adj_cols[s_adj[j] + p];
#pragma unroll
          for (int a = 0; a < RD; ++a)
#pragma unroll
            for (int b = 0; b < CD; ++b) indices[s_base[j * RD + a] + p * CD + b] = c * CD + b;
        }
      }
    }
    __syncthreads();
  }
}

// ---- fused integrate-and-assemble, interior nodes, affine elements --------------------------------------------------
// The element entries are EVALUATED -- from the element's coefficients (staged per batch in shared memory) and the
// product record of the local node pair (elem_affine.cuh; the table sits in shared memory when it fits, else it is read
// through L1) -- instead of being read from a materialised element matrix: the only HBM traffic left is the CSR rows
// written once, the row pointers, 1 B per (adjacency slot, incident element) of the inverse position map and 144 B of
// coefficients per element.
//
// A CTA takes ASM_NB consecutive nodes and its threads share the FLAT list of their adjacency slots (the nodes' slots
// are contiguous in the adjacency CSR, and so are their output rows): one RD x CD output block per thread and step,
// whatever the row lengths -- no per-node loop, no idle lanes on short rows.  A thread adds the contributions of the
// node's elements in ASCENDING cell order (the reference's cell loop: same order of additions per entry as
// asm_interior_kernel, hence the same bits).
constexpr int AFF_KC = 4;    // incident elements per node whose coefficients are staged (more: read from global)
constexpr int AFF_SLOTS = 4096;   // adjacency slots of a batch with a slot -> node table (longer batches: binary search)

template <int CD, bool TSM>
__global__ void __launch_bounds__(256, TSM ? 3 : 4)
asm_interior_affine_kernel(pyn::TopoView t, DescView d, const int64_t* __restrict__ indptr, int* __restrict__ indices,
                           double* __restrict__ data) {
  constexpr int RD = 2;
  constexpr int NC = CD == 2 ? 13 : 4, C0 = CD == 2 ? 0 : 13;     // coefficients used by this block shape
  constexpr int PR = CD == 2 ? pyn::PYN_AFF_PK : pyn::PYN_AFF_PR;
  extern __shared__ __align__(16) double s_tab[];                  // [nn*nn][PR] when TSM
  __shared__ int64_t s_inv[ASM_NB + 1], s_base[ASM_NB * RD];
  __shared__ int s_adj[ASM_NB + 1];                                // adjacency offsets relative to the batch
  __shared__ int s_el[ASM_NB * ASM_MAXDEG], s_lr[ASM_NB * ASM_MAXDEG];
  __shared__ __align__(16) uint8_t s_ivb[ASM_IVCAP + 32];
  __shared__ int s_n2e[ASM_NB + 1];
  __shared__ uint8_t s_flag[ASM_NB];
  __shared__ double s_cf[ASM_NB * AFF_KC * NC];
  __shared__ uint8_t s_q2j[AFF_SLOTS];                             // adjacency slot of the batch -> node of the batch
  const int tid = threadIdx.x;
  const int nn = t.nn;
  const double* gtab = CD == 2 ? d.aff.pairK : d.aff.pairR;
  if (TSM) {
    const int len = nn * nn * PR;
    for (int i = tid * 2; i < len; i += 512) *reinterpret_cast<double2*>(s_tab + i) = *reinterpret_cast<const double2*>(gtab + i);
  }
  const double* tab = TSM ? s_tab : gtab;
  const int64_t nbatch = (t.nnodes + ASM_NB - 1) / ASM_NB;
  for (int64_t batch = blockIdx.x; batch < nbatch; batch += gridDim.x) {
    const int64_t r0 = batch * ASM_NB;
    const int nb = (int)min((int64_t)ASM_NB, t.nnodes - r0);
    const int64_t e0 = t.n2e_ptr[r0], a0 = t.adj_ptr[r0];
    __syncthreads();
    if (tid <= nb) {
      s_adj[tid] = (int)(t.adj_ptr[r0 + tid] - a0);
      s_n2e[tid] = (int)(t.n2e_ptr[r0 + tid] - e0);
      s_inv[tid] = t.inv_ptr[r0 + tid];
    }
    if (tid < nb) s_flag[tid] = d.flags[r0 + tid];
    if (tid >= 64 && tid < 64 + nb * RD) s_base[tid - 64] = indptr[r0 * RD + (tid - 64)];
    __syncthreads();
    const uint8_t* ivsrc = t.inv + s_inv[0];
    const int mis = (int)(reinterpret_cast<uintptr_t>(ivsrc) & 15);
    const int64_t ivlen = s_inv[nb] - s_inv[0];
    const bool staged = ivlen + mis <= ASM_IVCAP;
    if (staged && data) {
      const uint4* src = reinterpret_cast<const uint4*>(ivsrc - mis);
      const int n16 = (int)((ivlen + mis + 15) >> 4);
      for (int i = tid; i < n16; i += 256) reinterpret_cast<uint4*>(s_ivb)[i] = src[i];
    }
    {  // (element, local row) of every (node, incident element): 8 threads per node
      const int j = tid >> 3, k = tid & 7;
      if (j < nb && !s_flag[j] && k < s_n2e[j + 1] - s_n2e[j]) {
        const int64_t q = e0 + s_n2e[j] + k;
        s_el[j * ASM_MAXDEG + k] = t.n2e_elem[q];
        s_lr[j * ASM_MAXDEG + k] = t.n2e_loc[q];
      }
    }
    __syncthreads();
    if (data) {  // coefficients of the first AFF_KC elements of every node: independent loads, issued back to back
      constexpr int PER = (ASM_NB * AFF_KC * NC + 255) / 256;
      double tmp[PER];
#pragma unroll
      for (int u = 0; u < PER; ++u) {
        const int i = tid + u * 256;
        const int j = i / (AFF_KC * NC), rem = i - j * (AFF_KC * NC), k = rem / NC, c = rem - k * NC;
        const bool ok = j < nb && !s_flag[j] && k < s_n2e[j + 1] - s_n2e[j];
        tmp[u] = ok ? __ldg(d.coef + (int64_t)s_el[j * ASM_MAXDEG + k] * PYN_AFF_NCOEF + C0 + c) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < PER; ++u) {
        const int i = tid + u * 256;
        if (i < ASM_NB * AFF_KC * NC) s_cf[i] = tmp[u];
      }
    }
    const int total = s_adj[nb];
    const bool table = total <= AFF_SLOTS;
    if (table) {   // slot -> node: 8 threads per node fill the node's range
      const int j = tid >> 3;
      if (j < nb)
        for (int q = s_adj[j] + (tid & 7); q < s_adj[j + 1]; q += 8) s_q2j[q] = (uint8_t)j;
    }
    __syncthreads();
    for (int q = tid; q < total; q += 256) {   // one output block per thread: adjacency slot q of the batch
      int j = 0;                               // its node: largest j with s_adj[j] <= q
      if (table) {
        j = s_q2j[q];
      } else {
#pragma unroll
        for (int step = ASM_NB / 2; step > 0; step >>= 1)
          if (j + step < nb && s_adj[j + step] <= q) j += step;
      }
      if (s_flag[j]) continue;
      const int p = q - s_adj[j], L = s_adj[j + 1] - s_adj[j], deg = s_n2e[j + 1] - s_n2e[j];
      if (data) {
        const uint8_t* ivj = staged ? s_ivb + mis + (int)(s_inv[j] - s_inv[0]) : t.inv + s_inv[j];
        double v[RD][CD];
#pragma unroll
        for (int a = 0; a < RD; ++a)
#pragma unroll
          for (int b = 0; b < CD; ++b) v[a][b] = 0.0;
        for (int k = 0; k < deg; ++k) {        // ascending cell order
          const int l = ivj[k * L + p];
          if (l == 0xFF) continue;
          const double* rp = tab + ((int64_t)s_lr[j * ASM_MAXDEG + k] * nn + l) * PR;
          double rec[PR];
#pragma unroll
          for (int i = 0; i < PR; i += 2) {
            const double2 x = TSM ? *reinterpret_cast<const double2*>(rp + i) : __ldg(reinterpret_cast<const double2*>(rp + i));
            rec[i] = x.x;
            rec[i + 1] = x.y;
          }
          double blk[RD][CD];
          if (k < AFF_KC) {
            affine_entry<RD, CD>(d, s_cf + (j * AFF_KC + k) * NC - C0, rec, blk);
          } else {
            affine_entry<RD, CD>(d, d.coef + (int64_t)s_el[j * ASM_MAXDEG + k] * PYN_AFF_NCOEF, rec, blk);
          }
#pragma unroll
          for (int a = 0; a < RD; ++a)
#pragma unroll
            for (int b = 0; b < CD; ++b) v[a][b] += d.sign * blk[a][b];
        }
#pragma unroll
        for (int a = 0; a < RD; ++a) {
          const int64_t base = s_base[j * RD + a];
          double* o = data + base + p * CD;
          if (CD == 2 && ((base & 1) == 0)) {
            *reinterpret_cast<double2*>(o) = make_double2(v[a][0], v[a][CD - 1]);
          } else {
#pragma unroll
            for (int b = 0; b < CD; ++b) o[b] = v[a][b];
          }
        }
      }
      if (indices) {
        const int c = t.adj_cols[a0 + q];
#pragma unroll
        for (int a = 0; a < RD; ++a)
#pragma unroll
          for (int b = 0; b < CD; ++b) indices[s_base[j * RD + a] + p * CD + b] = c * CD + b;
      }
    }
  }
}

// ---- fused integrate-and-assemble on a STRUCTURED mesh: nodes grouped by row signature ---------------------------------
// The flat kernel above spends most of its instructions on finding out, per output block, which node it belongs to and
// which of the node's elements contribute (and its element loop diverges).  On a structured mesh that information is the
// same for every node of a row signature (topo.cuh: byte-identical incidence local indices and inverse maps).  Here a
// WARP takes a tile of 32 nodes of ONE signature and walks their rows in passes:
//   * a row of nut units (a unit = one double2 of each of the node's two rows = 1 adjacency position of a 2x2-block
//     matrix, 2 positions of a 2x1-block one) is cut into passes whose widths are the powers of two of nut's binary
//     expansion, largest first, at most 32: 2^s lanes share a node, one unit per lane, 32 >> s nodes per step, 2^s
//     steps per pass.  Every lane is busy whatever the row length and one store instruction writes pieces as long as
//     the row allows (512 contiguous bytes when 2^s = 32);
//   * a lane's unit is fixed during a pass, so the product records it needs (one per incident element: the
//     representative node's inverse map says which) sit in its REGISTERS for all 32 nodes -- the inverse-map bytes are
//     fetched two passes ahead and, where the registers allow (valence 1), the records one pass ahead;
//   * what changes from step to step is the node: the 13 (4) coefficients of its <= DEG elements are staged once per
//     tile in shared memory (cp.async) and read as broadcast LDS.128;
//   * the lanes of a node store consecutive double2 of each of its two rows straight from registers: no output tile,
//     no transposition, every step independent of the others.
// Per tile the kernel reads a 16-byte descriptor and one slot per lane {node, its elements} (topo.cuh), so two dependent
// loads precede the first pass.  DEG (incident elements per node) is a template parameter: the tiles are sorted by
// valence class and each class gets its own launch.  Same order of additions per entry as every other assembly kernel
// (ascending cell order): same bits.
constexpr int SIG_W = 4;       // warps (= tiles) per CTA
constexpr int SIG_KMAX = 4;    // incident elements per node (quadrilateral meshes; larger valence: flat kernel)

// v2 store when the address is 16-byte aligned, else one or two scalar stores -- predicated, no branches
__device__ __forceinline__ void sig_store(double* o, double x0, double x1, bool v2, bool lo, bool hi) {
  asm volatile(
      "{\n\t.reg .pred p2, pl, ph;\n\t"
      "setp.ne.b32 p2, %3, 0;\n\tsetp.ne.b32 pl, %4, 0;\n\tsetp.ne.b32 ph, %5, 0;\n\t"
      "@p2 st.global.v2.f64 [%0], {%1, %2};\n\t"
      "@pl st.global.f64 [%0], %1;\n\t"
      "@ph st.global.f64 [%0+8], %2;\n\t}"
      :
      : "l"(o), "d"(x0), "d"(x1), "r"((int)v2), "r"((int)lo), "r"((int)hi)
      : "memory");
}
template <int BYTES>
__device__ __forceinline__ void sig_cp_async(void* smem_dst, const void* gsrc) {
  const unsigned dst = (unsigned)__cvta_generic_to_shared(smem_dst);
  if constexpr (BYTES == 16) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(gsrc) : "memory");
  } else {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(dst), "l"(gsrc) : "memory");
  }
}
__device__ __forceinline__ void sig_cp_async_wait() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }

template <int CD, int DEG>
struct SigShape {
  static constexpr int RD = 2;
  static constexpr int NP = 2 / CD;                                  // adjacency positions per unit
  static constexpr int PR = CD == 2 ? pyn::PYN_AFF_PK : pyn::PYN_AFF_PR;
  static constexpr int NC = CD == 2 ? 13 : 4, NCP = CD == 2 ? 14 : 4;   // coefficients per element (padded to 16 B)
  static constexpr int SL = DEG == 1 ? 2 : (DEG == 2 ? 4 : 8);       // ints per node slot
  static constexpr int per_warp = 32 * DEG * NCP + 32 * RD;          // 8-byte words of shared memory per warp
  static constexpr bool DBL = DEG == 1;                              // records double-buffered in registers
};

template <int CD, int DEG>
__global__ void __launch_bounds__(SIG_W * 32, DEG == 4 ? 3 : 4)
asm_sig_affine_kernel(pyn::TopoView t, DescView d, const int64_t* __restrict__ indptr, double* __restrict__ data,
                      int64_t tile0, int64_t tile1, const int* __restrict__ slots) {
  using S = SigShape<CD, DEG>;
  constexpr int RD = 2, PR = S::PR, NP = S::NP, NC = S::NC, NCP = S::NCP, SL = S::SL;
  constexpr int C0 = CD == 2 ? 0 : 13;
  constexpr bool DBL = S::DBL;
  extern __shared__ __align__(16) double sg_dyn[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double* s_cf = sg_dyn + (size_t)w * S::per_warp;                   // [node][k][NCP]
  double** s_ptr = reinterpret_cast<double**>(s_cf + 32 * DEG * NCP);   // [node][a]: start of row a of the node
  const int64_t tile = tile0 + (int64_t)blockIdx.x * SIG_W + w;
  if (tile >= tile1) return;                                         // warps are independent: no CTA-wide barrier below
  const double* gtab = CD == 2 ? d.aff.pairK : d.aff.pairR;
  const int nn = t.nn;
  const double sign = d.sign;
  // level 1: the tile's descriptor and this lane's node slot
  const int4 dd = __ldg(reinterpret_cast<const int4*>(d.tile_desc) + tile);
  int sl[1 + DEG];
  {
    const int* sp = slots + ((tile - tile0) * 32 + lane) * SL;
    if constexpr (SL == 2) {
      const int2 x = __ldg(reinterpret_cast<const int2*>(sp));
      sl[0] = x.x; sl[1] = x.y;
    } else {
      const int4 x = __ldg(reinterpret_cast<const int4*>(sp));
      sl[0] = x.x; sl[1] = x.y; sl[2] = x.z;
      if constexpr (SL == 8) {
        sl[3] = x.w;
        sl[4] = __ldg(sp + 4);
      }
    }
  }
  const int cnt = dd.x & 255, deg = min((dd.x >> 8) & 255, DEG), L = (int)((unsigned)dd.x >> 16);
  const unsigned lrp = (unsigned)dd.y;
  const uint8_t* iv = t.inv + (((int64_t)(dd.w & 0xffffff) << 32) | (int64_t)(unsigned)dd.z);   // (top byte: row offsets)
  const int nut = (L + NP - 1) / NP;                                 // units of a row
  auto width_log2 = [&](int left) { return min(5, 31 - __clz(left)); };   // left >= 1 units remain
  // inverse-map bytes of this lane's unit in the pass that starts at unit pu (0xFF: no contribution / no position)
  auto load_l = [&](int pu, int (&l)[DEG][NP]) {
    const int s = pu < nut ? width_log2(nut - pu) : 0;
    const int u = lane & ((1 << s) - 1);
#pragma unroll
    for (int k = 0; k < DEG; ++k)
#pragma unroll
      for (int h = 0; h < NP; ++h) {
        const int q = (pu + u) * NP + h;
        l[k][h] = (pu < nut && k < deg && q < L) ? (int)iv[k * L + q] : 0xFF;
      }
  };
  auto load_rec = [&](const int (&l)[DEG][NP], double (&rec)[DEG][NP][PR]) {
#pragma unroll
    for (int k = 0; k < DEG; ++k)
#pragma unroll
      for (int h = 0; h < NP; ++h)
        if (l[k][h] != 0xFF) {
          const double2* rp =
              reinterpret_cast<const double2*>(gtab + ((int64_t)((lrp >> (8 * k)) & 255u) * nn + l[k][h]) * PR);
#pragma unroll
          for (int i = 0; i < PR / 2; ++i) {
            const double2 x = __ldg(rp + i);
            rec[k][h][2 * i] = x.x;
            rec[k][h][2 * i + 1] = x.y;
          }
        }
  };
  // level 2: coefficients of the lane's node -> shared memory (asynchronous), its row starts, the inverse-map bytes of
  // the first two passes
#pragma unroll
  for (int k = 0; k < DEG; ++k) {
    const double* cg = d.coef + (int64_t)sl[1 + k] * PYN_AFF_NCOEF + C0;
    double* dst = s_cf + (lane * DEG + k) * NCP;
    if constexpr (CD == 2) {
#pragma unroll
      for (int i = 0; i < NCP / 2; ++i) sig_cp_async<16>(dst + 2 * i, cg + 2 * i);
    } else {
#pragma unroll
      for (int i = 0; i < NC; ++i) sig_cp_async<8>(dst + i, cg + i);
    }
  }
  const int64_t q0 = indptr[(int64_t)sl[0] * RD], q1 = indptr[(int64_t)sl[0] * RD + 1];
  int l0[DEG][NP], l1[DEG][NP];
  load_l(0, l0);
  load_l(1 << width_log2(nut), l1);
  double rec[DEG][NP][PR], recn[DBL ? DEG : 1][NP][PR];
#pragma unroll
  for (int k = 0; k < DEG; ++k)
#pragma unroll
    for (int h = 0; h < NP; ++h)
#pragma unroll
      for (int i = 0; i < PR; ++i) {
        rec[k][h][i] = 0.0;
        if (DBL) recn[k][h][i] = 0.0;
      }
  s_ptr[lane * RD] = data + q0;
  s_ptr[lane * RD + 1] = data + q1;
  load_rec(l0, rec);                                                 // level 3
  sig_cp_async_wait();
  __syncwarp();

  int sn = width_log2(nut);
  for (int pu = 0; pu < nut;) {
    const int s = sn;
    const int u = lane & ((1 << s) - 1), grp = lane >> s, npg = 32 >> s;
    const int pu1 = pu + (1 << s);                           // the next pass, and the one after
    sn = pu1 < nut ? width_log2(nut - pu1) : 0;
    const int pu2 = pu1 + (1 << sn);
    if constexpr (DBL) load_rec(l1, recn);
    int l2[DEG][NP];
    load_l(pu2, l2);
    bool on[DEG][NP], anyk[DEG];
#pragma unroll
    for (int k = 0; k < DEG; ++k) {
      bool some = false;
#pragma unroll
      for (int h = 0; h < NP; ++h) {
        on[k][h] = l0[k][h] != 0xFF;
        some = some || on[k][h];
      }
      anyk[k] = __any_sync(0xffffffffu, some);
    }
    const bool second = CD == 2 || (pu + u) * NP + 1 < L;    // the unit's second double exists
    const int64_t off = (int64_t)(pu + u) * 2;
#pragma unroll 2
    for (int st = 0; st < (1 << s); ++st) {                  // npg nodes per step
      const int n = st * npg + grp;
      const double* cfn = s_cf + n * (DEG * NCP);
      const longlong2 pp = *reinterpret_cast<const longlong2*>(s_ptr + n * RD);
      double v[NP][RD][CD];
#pragma unroll
      for (int h = 0; h < NP; ++h)
#pragma unroll
        for (int a = 0; a < RD; ++a)
#pragma unroll
          for (int b = 0; b < CD; ++b) v[h][a][b] = 0.0;
#pragma unroll
      for (int k = 0; k < DEG; ++k) {                        // ascending cell order
        if (DEG == 1 || anyk[k]) {                           // uniform: some lane of the warp takes element k
          double cf[NCP];
#pragma unroll
          for (int i = 0; i < NCP / 2; ++i) {
            const double2 x = *reinterpret_cast<const double2*>(cfn + k * NCP + 2 * i);
            cf[2 * i] = x.x;
            cf[2 * i + 1] = x.y;
          }
#pragma unroll
          for (int h = 0; h < NP; ++h) {
            double blk[RD][CD];
            if constexpr (CD == 2) {
              pyn::affine_block_K(cf, rec[k][h], blk);
            } else {
              pyn::affine_block_R(cf, rec[k][h], d.which, d.aff.alpha_w, d.aff.alpha_d, blk);
            }
            if (on[k][h]) {
#pragma unroll
              for (int a = 0; a < RD; ++a)
#pragma unroll
                for (int b = 0; b < CD; ++b) v[h][a][b] += sign * blk[a][b];
            }
          }
        }
      }
      const bool ok = n < cnt;
#pragma unroll
      for (int a = 0; a < RD; ++a) {
        double* o = reinterpret_cast<double*>(a == 0 ? pp.x : pp.y) + off;
        const bool al = ok && second && (reinterpret_cast<uintptr_t>(o) & 15) == 0;
        sig_store(o, v[0][a][0], CD == 2 ? v[0][a][CD - 1] : v[NP - 1][a][0], al, ok && !al, ok && !al && second);
      }
    }
    // rotate the look-ahead
#pragma unroll
    for (int k = 0; k < DEG; ++k)
#pragma unroll
      for (int h = 0; h < NP; ++h) {
        l0[k][h] = l1[k][h];
        l1[k][h] = l2[k][h];
        if constexpr (DBL) {
#pragma unroll
          for (int i = 0; i < PR; ++i) rec[k][h][i] = recn[k][h][i];
        }
      }
    if constexpr (!DBL) load_rec(l0, rec);
    pu = pu1;
  }
}

// ---- the same with LINE-ALIGNED stores ---------------------------------------------------------------------------------
// The rows of a CSR matrix start wherever the previous row ended: 16 bytes into a 32-byte sector every second time when a
// row holds an odd number of 2x2 blocks, 8 bytes off for 2x1 blocks.  Measured on the kernel above (pieces rounded down
// to 128-byte lines, wrong addresses): the misaligned pieces cost 20 % (K) to 35 % (Rw) of the launch.  Here the tiles
// are grouped by (signature, offset of the node's first row inside a 128-byte line) -- ensure_sig_tiles sorts by both, the
// second row's offset follows from the row length -- and a lane's pair of doubles is fixed in LINE coordinates: pair y
// of row a is the row's double y - tau_a, so the 2^s lanes of a node always cover whole, aligned 16-byte units and a
// 32-lane pass writes four whole lines.  The price: the two rows of a node are evaluated at DIFFERENT column positions
// (tau_0 != tau_1), each from its own product record (twice the record registers) with the row-wise closed form
// (affine_row_K / affine_row_R: the same operations per entry as the block form, so the same bits).
template <int CD, int DEG>
struct SigRowShape {
  static constexpr int NP = 2 / CD;                                  // adjacency positions per pair of doubles
  static constexpr int PR = CD == 2 ? pyn::PYN_AFF_PK : pyn::PYN_AFF_PR;
  static constexpr int NC = CD == 2 ? 13 : 4, NCP = CD == 2 ? 14 : 4;
  static constexpr int SL = DEG == 1 ? 2 : (DEG == 2 ? 4 : 8);
  static constexpr int per_warp = 32 * DEG * NCP + 32 * 2;           // 8-byte words of shared memory per warp
};

template <int CD, int DEG>
__global__ void __launch_bounds__(SIG_W * 32, DEG == 1 ? 4 : 3)
asm_sigrow_affine_kernel(pyn::TopoView t, DescView d, const int64_t* __restrict__ indptr, double* __restrict__ data,
                         int64_t tile0, int64_t tile1, const int* __restrict__ slots) {
  using S = SigRowShape<CD, DEG>;
  constexpr int RD = 2, PR = S::PR, NP = S::NP, NC = S::NC, NCP = S::NCP, SL = S::SL;
  constexpr int C0 = CD == 2 ? 0 : 13;
  extern __shared__ __align__(16) double sg_dyn[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  double* s_cf = sg_dyn + (size_t)w * S::per_warp;                   // [node][k][NCP]
  double** s_ptr = reinterpret_cast<double**>(s_cf + 32 * DEG * NCP);   // [node][a]: start of row a of the node
  const int64_t tile = tile0 + (int64_t)blockIdx.x * SIG_W + w;
  if (tile >= tile1) return;
  const double* gtab = CD == 2 ? d.aff.pairK : d.aff.pairR;
  const int nn = t.nn;
  const double sign = d.sign;
  const int4 dd = __ldg(reinterpret_cast<const int4*>(d.tile_desc) + tile);
  int sl[1 + DEG];
  {
    const int* sp = slots + ((tile - tile0) * 32 + lane) * SL;
    if constexpr (SL == 2) {
      const int2 x = __ldg(reinterpret_cast<const int2*>(sp));
      sl[0] = x.x; sl[1] = x.y;
    } else {
      const int4 x = __ldg(reinterpret_cast<const int4*>(sp));
      sl[0] = x.x; sl[1] = x.y; sl[2] = x.z;
      if constexpr (SL == 8) {
        sl[3] = x.w;
        sl[4] = __ldg(sp + 4);
      }
    }
  }
  const int cnt = dd.x & 255, deg = min((dd.x >> 8) & 255, DEG), L = (int)((unsigned)dd.x >> 16);
  const unsigned lrp = (unsigned)dd.y;
  const uint8_t* iv = t.inv + (((int64_t)(dd.w & 0xffffff) << 32) | (int64_t)(unsigned)dd.z);
  // offsets of the two rows inside a line, in doubles (2x2 blocks: rounded down to a whole block -- a row that starts on
  // an odd double keeps its scalar stores)
  int tau[RD] = {(dd.w >> 24) & 15, (int)((unsigned)dd.w >> 28)};
  if (CD == 2) {
    tau[0] &= ~1;
    tau[1] &= ~1;
  }
  const int rowlen = L * CD;
  const int nuy = (max(tau[0], tau[1]) + rowlen + 1) / 2;            // pairs of a row in line coordinates
  auto width_log2 = [&](int left) { return min(5, 31 - __clz(left)); };
  auto load_l = [&](int pu, int (&l)[RD][DEG][NP]) {
    const int s = pu < nuy ? width_log2(nuy - pu) : 0;
    const int u = lane & ((1 << s) - 1);
#pragma unroll
    for (int a = 0; a < RD; ++a) {
      const int tt = 2 * (pu + u) - tau[a];                          // row-relative double of the pair's first entry
#pragma unroll
      for (int k = 0; k < DEG; ++k)
#pragma unroll
        for (int h = 0; h < NP; ++h) {
          const int q = CD == 2 ? (tt >> 1) : tt + h;
          const bool ok = pu < nuy && k < deg && (CD == 2 ? tt >= 0 : q >= 0) && q < L;
          l[a][k][h] = ok ? (int)iv[k * L + q] : 0xFF;
        }
    }
  };
  auto load_rec = [&](const int (&l)[RD][DEG][NP], double (&rec)[RD][DEG][NP][PR]) {
#pragma unroll
    for (int a = 0; a < RD; ++a)
#pragma unroll
      for (int k = 0; k < DEG; ++k)
#pragma unroll
        for (int h = 0; h < NP; ++h)
          if (l[a][k][h] != 0xFF) {
            const double2* rp =
                reinterpret_cast<const double2*>(gtab + ((int64_t)((lrp >> (8 * k)) & 255u) * nn + l[a][k][h]) * PR);
#pragma unroll
            for (int i = 0; i < PR / 2; ++i) {
              const double2 x = __ldg(rp + i);
              rec[a][k][h][2 * i] = x.x;
              rec[a][k][h][2 * i + 1] = x.y;
            }
          }
  };
#pragma unroll
  for (int k = 0; k < DEG; ++k) {
    const double* cg = d.coef + (int64_t)sl[1 + k] * PYN_AFF_NCOEF + C0;
    double* dst = s_cf + (lane * DEG + k) * NCP;
    if constexpr (CD == 2) {
#pragma unroll
      for (int i = 0; i < NCP / 2; ++i) sig_cp_async<16>(dst + 2 * i, cg + 2 * i);
    } else {
#pragma unroll
      for (int i = 0; i < NC; ++i) sig_cp_async<8>(dst + i, cg + i);
    }
  }
  const int64_t q0 = indptr[(int64_t)sl[0] * RD], q1 = indptr[(int64_t)sl[0] * RD + 1];
  int l0[RD][DEG][NP], l1[RD][DEG][NP];
  load_l(0, l0);
  load_l(1 << width_log2(nuy), l1);
  double rec[RD][DEG][NP][PR];
#pragma unroll
  for (int a = 0; a < RD; ++a)
#pragma unroll
    for (int k = 0; k < DEG; ++k)
#pragma unroll
      for (int h = 0; h < NP; ++h)
#pragma unroll
        for (int i = 0; i < PR; ++i) rec[a][k][h][i] = 0.0;
  // row starts in line coordinates: pair y of row a sits at s_ptr[a] + 2 y
  s_ptr[lane * RD] = data + q0 - tau[0];
  s_ptr[lane * RD + 1] = data + q1 - tau[1];
  load_rec(l0, rec);
  sig_cp_async_wait();
  __syncwarp();

  int sn = width_log2(nuy);
  for (int pu = 0; pu < nuy;) {
    const int s = sn;
    const int u = lane & ((1 << s) - 1), grp = lane >> s, npg = 32 >> s;
    const int pu1 = pu + (1 << s);
    sn = pu1 < nuy ? width_log2(nuy - pu1) : 0;
    int l2[RD][DEG][NP];
    load_l(pu1 + (1 << sn), l2);
    bool on[RD][DEG][NP], anyk[DEG], first[RD], second[RD];
#pragma unroll
    for (int k = 0; k < DEG; ++k) {
      bool some = false;
#pragma unroll
      for (int a = 0; a < RD; ++a)
#pragma unroll
        for (int h = 0; h < NP; ++h) {
          on[a][k][h] = l0[a][k][h] != 0xFF;
          some = some || on[a][k][h];
        }
      anyk[k] = __any_sync(0xffffffffu, some);
    }
#pragma unroll
    for (int a = 0; a < RD; ++a) {                           // which doubles of the pair exist in row a
      const int tt = 2 * (pu + u) - tau[a];
      first[a] = tt >= 0 && tt < rowlen;
      second[a] = tt + 1 >= 0 && tt + 1 < rowlen;
    }
    const int64_t off = (int64_t)(pu + u) * 2;
#pragma unroll 2
    for (int st = 0; st < (1 << s); ++st) {
      const int n = st * npg + grp;
      const double* cfn = s_cf + n * (DEG * NCP);
      const longlong2 pp = *reinterpret_cast<const longlong2*>(s_ptr + n * RD);
      double v[RD][2];
#pragma unroll
      for (int a = 0; a < RD; ++a) v[a][0] = v[a][1] = 0.0;
#pragma unroll
      for (int k = 0; k < DEG; ++k) {                        // ascending cell order
        if (DEG == 1 || anyk[k]) {
          double cf[NCP];
#pragma unroll
          for (int i = 0; i < NCP / 2; ++i) {
            const double2 x = *reinterpret_cast<const double2*>(cfn + k * NCP + 2 * i);
            cf[2 * i] = x.x;
            cf[2 * i + 1] = x.y;
          }
#pragma unroll
          for (int a = 0; a < RD; ++a) {
            if constexpr (CD == 2) {
              double r2[2];
              pyn::affine_row_K(cf, rec[a][k][0], a, r2);
              if (on[a][k][0]) {
                v[a][0] += sign * r2[0];
                v[a][1] += sign * r2[1];
              }
            } else {
#pragma unroll
              for (int h = 0; h < NP; ++h) {
                const double r1 = pyn::affine_row_R(cf, rec[a][k][h], d.which, d.aff.alpha_w, d.aff.alpha_d, a);
                if (on[a][k][h]) v[a][h] += sign * r1;
              }
            }
          }
        }
      }
      const bool okn = n < cnt;
#pragma unroll
      for (int a = 0; a < RD; ++a) {
        double* o = reinterpret_cast<double*>(a == 0 ? pp.x : pp.y) + off;
        const bool f = okn && first[a], g2 = okn && second[a];
        const bool al = f && g2 && (reinterpret_cast<uintptr_t>(o) & 15) == 0;
        sig_store(o, v[a][0], v[a][1], al, f && !al, g2 && !al);
      }
    }
#pragma unroll
    for (int a = 0; a < RD; ++a)
#pragma unroll
      for (int k = 0; k < DEG; ++k)
#pragma unroll
        for (int h = 0; h < NP; ++h) {
          l0[a][k][h] = l1[a][k][h];
          l1[a][k][h] = l2[a][k][h];
        }
    load_rec(l0, rec);
    pu = pu1;
  }
}

// ---- uniform-mesh fast path: one template per row signature, streamed to every node that carries it ----------------
// With ONE element matrix for all cells (E_stride == 0: the reference integrates cell 0 only, base_problem.py:333,133)
// the rows of an interior node depend only on its signature (topo.cuh).  asm_template_kernel computes the RD x (L*CD)
// block of each signature with exactly the arithmetic of asm_interior_kernel (same order of additions -> same bits);
// asm_stream_kernel then writes it to every unflagged node: a pure store stream, bounded by HBM write bandwidth.
template <int RD, int CD>
__global__ void __launch_bounds__(128)
asm_template_kernel(pyn::TopoView t, DescView d, double* __restrict__ T) {
  const int lane = threadIdx.x & 31;
  const int ldE = CD * t.nn;
  for (int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < t.nsig; s += gridDim.x * (blockDim.x >> 5)) {
    const int64_t r = t.sig_rep[s];
    const int64_t b0 = t.n2e_ptr[r];
    const int L = (int)(t.adj_ptr[r + 1] - t.adj_ptr[r]), deg = (int)(t.n2e_ptr[r + 1] - b0);
    const uint8_t* iv = t.inv + t.inv_ptr[r];
    double* out = T + t.sig_ptr[s] * (RD * CD);
    for (int p = lane; p < L; p += 32) {
      double v[RD][CD];
#pragma unroll
      for (int a = 0; a < RD; ++a)
#pragma unroll
        for (int b = 0; b < CD; ++b) v[a][b] = 0.0;
      for (int k = 0; k < deg; ++k) {  // ascending cell order
        const int l = iv[k * L + p];
        if (l != 0xFF) {
          const double* Ee = d.E + (int64_t)t.n2e_loc[b0 + k] * RD * ldE + l * CD;
#pragma unroll
          for (int a = 0; a < RD; ++a)
#pragma unroll
            for (int b = 0; b < CD; ++b) v[a][b] += d.sign * Ee[a * ldE + b];
        }
      }
#pragma unroll
      for (int a = 0; a < RD; ++a)
#pragma unroll
        for (int b = 0; b < CD; ++b) out[(a * L + p) * CD + b] = v[a][b];
    }
  }
}

__device__ int g_stream_warp_only = 0;   // PYN_ASM_WARP_ONLY=1: keep the per-warp streaming only (A/B measurements)
constexpr int ASM_SB = 256;  // nodes per CTA batch of the stream kernel
constexpr int ASM_SG = ASM_SB / 8;   // consecutive nodes per warp
template <int RD, int CD>
__global__ void __launch_bounds__(256)
asm_stream_kernel(pyn::TopoView t, DescView d, const int64_t* __restrict__ indptr, const double* __restrict__ T,
                  double* __restrict__ data) {
  __shared__ int64_t s_base[ASM_SB], s_src[ASM_SB];
  __shared__ int s_len[ASM_SB];
  __shared__ double s_rs[ASM_SB];
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int64_t nbatch = (t.nnodes + ASM_SB - 1) / ASM_SB;
  for (int64_t batch = blockIdx.x; batch < nbatch; batch += gridDim.x) {
    const int64_t r0 = batch * ASM_SB;
    const int nb = (int)min((int64_t)ASM_SB, t.nnodes - r0);
    if (tid < nb) {
      const int64_t r = r0 + tid;
      const bool skip = d.flags[r] != 0;
      const int L = (int)(t.adj_ptr[r + 1] - t.adj_ptr[r]);
      s_len[tid] = skip ? 0 : L * (RD * CD);        // the RD rows of an interior node are contiguous in the CSR data
      s_base[tid] = indptr[r * RD];
      s_src[tid] = t.sig_ptr[t.sig_id[r]] * (RD * CD);
      s_rs[tid] = d.rowscale ? d.rowscale[r] : 1.0;
    }
    __syncthreads();
    // a warp takes ASM_SG consecutive nodes.  When their rows are equally long, adjacent in the CSR data and 16-byte
    // aligned (runs of cell-interior or edge nodes: almost always), the group is ONE contiguous destination range:
    // full-warp double2 stores, the source template found by a division -- short rows no longer leave lanes idle.
    // Whole batch uniform (the common case away from the walls and from the borders between node kinds): the CTA
    // streams the batch's one contiguous destination range with all 256 threads, so that at any instant its 8 warps
    // store 4 kB of ADJACENT addresses -- the access order of a plain fill, which is what HBM write-back likes.
    bool oka = true;
    if (tid < nb) {
      const int n00 = s_len[0];
      oka = n00 > 0 && s_len[tid] == n00 && s_base[tid] == s_base[0] + (int64_t)tid * n00 &&
            ((s_src[tid] | s_base[0] | n00) & 1) == 0;
    }
    const bool cta_uniform = __syncthreads_and(oka) && !g_stream_warp_only;
    if (cta_uniform) {
      const int n2 = s_len[0] >> 1, total2 = nb * n2;
      const unsigned magic = (unsigned)((0x100000000ull + (unsigned)n2 - 1) / (unsigned)n2);
      const double2* T2 = reinterpret_cast<const double2*>(T);
      double2* d2 = reinterpret_cast<double2*>(data + s_base[0]);
#pragma unroll 4
      for (int i = tid; i < total2; i += ASM_SB) {
        const int jj = (int)__umulhi((unsigned)i, magic), k = i - jj * n2;
        double2 v = __ldg(T2 + (s_src[jj] >> 1) + k);
        if (d.rowscale) { const double rs = s_rs[jj]; v.x = rs * v.x; v.y = rs * v.y; }
        d2[i] = v;
      }
      __syncthreads();
      continue;
    }
    const int j0 = w * ASM_SG, cnt = min(ASM_SG, nb - j0);
    if (cnt > 0) {
      const int n0 = s_len[j0];
      bool okl = true;
      if (lane < cnt) {
        const int j = j0 + lane;
        okl = s_len[j] == n0 && s_base[j] == s_base[j0] + (int64_t)lane * n0 && ((s_src[j] | s_base[j0] | n0) & 1) == 0;
      }
      const bool uniform = n0 > 0 && __all_sync(0xffffffffu, okl);
      if (uniform) {
        const int n2 = n0 >> 1, total2 = cnt * n2;
        const unsigned magic = (unsigned)((0x100000000ull + (unsigned)n2 - 1) / (unsigned)n2);
        const double2* T2 = reinterpret_cast<const double2*>(T);
        double2* d2 = reinterpret_cast<double2*>(data + s_base[j0]);
        for (int i = lane; i < total2; i += 32) {
          const int jj = (int)__umulhi((unsigned)i, magic), k = i - jj * n2;
          double2 v = __ldg(T2 + (s_src[j0 + jj] >> 1) + k);
          if (d.rowscale) { const double rs = s_rs[j0 + jj]; v.x = rs * v.x; v.y = rs * v.y; }
          d2[i] = v;
        }
      } else {
        for (int j = j0; j < j0 + cnt; ++j) {
          const int n = s_len[j];
          if (n == 0) continue;
          const double* src = T + s_src[j];
          double* dst = data + s_base[j];
          const double rs = s_rs[j];
          if (((s_base[j] | s_src[j] | n) & 1) == 0) {      // 16-byte aligned on both sides: double2 stores
            const double2* s2 = reinterpret_cast<const double2*>(src);
            double2* d2 = reinterpret_cast<double2*>(dst);
            for (int i = lane; i < (n >> 1); i += 32) {
              double2 v = __ldg(s2 + i);
              if (d.rowscale) { v.x = rs * v.x; v.y = rs * v.y; }
              d2[i] = v;
            }
          } else {
            for (int i = lane; i < n; i += 32) {
              double v = __ldg(src + i);
              if (d.rowscale) v = rs * v;
              dst[i] = v;
            }
          }
        }
      }
    }
    __syncthreads();
  }
}

// row lengths of the interior nodes: L*cdof when the (free,free) pair is kept, else 0
__global__ void asm_interior_count_kernel(pyn::TopoView t, const uint8_t* __restrict__ flags, int rdof, int cdof,
                                          int keep00, int* __restrict__ rowcnt) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < t.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    if (flags[r]) continue;
    const int n = keep00 ? (int)(t.adj_ptr[r + 1] - t.adj_ptr[r]) * cdof : 0;
    for (int a = 0; a < rdof; ++a) rowcnt[r * rdof + a] = n;
  }
}

// A device scalar is handed to the host through mapped pinned memory (a store from a one-thread kernel), not through
// a device->host memcpy: a memcpy would queue on the copy engine behind whatever bulk download another stream has in
// flight (bench.py's end-to-end step downloads 268 MB on a second stream) and stall this stream for milliseconds.
__global__ void publish_i64_kernel(const int64_t* __restrict__ src, volatile int64_t* dst) { *dst = *src; }
int read_device_i64(const int64_t* d_src, int64_t* out, cudaStream_t st) {
  static int64_t* h_box = nullptr;     // one host thread per GPU (include/pynama_b200.h)
  static int64_t* d_box = nullptr;
  if (!h_box) {
    PYN_CHECK_CUDA(cudaHostAlloc(&h_box, sizeof(int64_t), cudaHostAllocMapped));
    PYN_CHECK_CUDA(cudaHostGetDevicePointer(&d_box, h_box, 0));
  }
  publish_i64_kernel<<<1, 1, 0, st>>>(d_src, d_box);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  *out = *h_box;
  return PYN_OK;
}

// bcmap: flag = node has a non-zero row class, or any adjacency column has a non-zero column class.  The node
// adjacency is symmetric, so the second condition is scattered from the (few) nodes that carry a non-zero column
// class instead of being searched from every node: O(nodes + boundary * row length) instead of O(adjacency).
__global__ void bcmap_own_kernel(pyn::TopoView t, const int8_t* __restrict__ rclass, int rdof, uint8_t* __restrict__ flags) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < t.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    bool f = (t.n2e_ptr[r + 1] - t.n2e_ptr[r]) > ASM_MAXDEG;  // high-valence nodes take the generic kernel
    if (rclass)
      for (int a = 0; a < rdof; ++a) f = f || (rclass[r * rdof + a] != 0);
    flags[r] = f ? 1 : 0;
  }
}
__global__ void bcmap_scatter_kernel(pyn::TopoView t, const int8_t* __restrict__ cclass, int cdof,
                                     uint8_t* __restrict__ flags) {
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < t.nnodes; c += (int64_t)gridDim.x * blockDim.x) {
    bool nz = false;
    for (int b = 0; b < cdof; ++b) nz = nz || (cclass[c * cdof + b] != 0);
    if (!nz) continue;
    for (int64_t q = t.adj_ptr[c]; q < t.adj_ptr[c + 1]; ++q) flags[t.adj_cols[q]] = 1;   // all writers store 1
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

__global__ void csr_diag_offsets_kernel(int64_t nrows, const int64_t* __restrict__ p, const int* __restrict__ ind,
                                        int* __restrict__ off) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = p[r], e = p[r + 1];
    int64_t lo = b, hi = e;
    while (lo < hi) {
      const int64_t mid = (lo + hi) >> 1;
      if (ind[mid] < r) lo = mid + 1; else hi = mid;
    }
    off[r] = (lo < e && ind[lo] == r) ? (int)(lo - b) : -1;
  }
}
__global__ void csr_diag_at_kernel(int64_t nrows, const int64_t* __restrict__ p, const int* __restrict__ off,
                                   const double* __restrict__ val, double* __restrict__ diag) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nrows; r += (int64_t)gridDim.x * blockDim.x) {
    const int o = off[r];
    diag[r] = o >= 0 ? val[p[r] + o] : 0.0;
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
  v->flags = desc->bcmap ? desc->bcmap->flags : nullptr;
  v->coef = nullptr; v->which = 0;
  v->tile_desc = v->tile_slot = nullptr; v->ntiles = 0;
  memset(&v->aff, 0, sizeof(v->aff));
  if (desc->d_coef) {   // fused integrate-and-assemble on affine elements
    PYN_REQUIRE(desc->elem != nullptr, "affine assembly: element handle missing");
    PYN_REQUIRE(desc->rdof == 2 && ((desc->which == 0 && desc->cdof == 2) || ((desc->which == 1 || desc->which == 2) && desc->cdof == 1)),
                "affine assembly: K (2x2) or Rw / Rd (2x1) blocks only");
    PYN_REQUIRE(desc->d_rowscale == nullptr, "affine assembly: no row scaling");
    v->coef = desc->d_coef; v->which = desc->which; v->aff = pyn::aff_tables(desc->elem);
    v->E = nullptr; v->E_stride = 1;   // per-cell by nature: the row-template shortcut does not apply
  }
  return PYN_OK;
}

}  // namespace

extern "C" {

}  // extern "C"

template <bool COUNT>
static int launch_asm(const pyn_topo* t, const DescView& dv, const int* list, int64_t nlist, int* cnt,
                      const int64_t* indptr, int* indices, double* data, cudaStream_t st) {
  const pyn::TopoView tv = pyn::view_of(t);
  const int64_t nwork = list ? nlist : t->nnodes;
  if (nwork == 0) return PYN_OK;
  const int grid = pyn::grid_for(nwork * 32, 256, 16);
  if (dv.coef && !COUNT) {
    PYN_REQUIRE(tv.nn == dv.aff.n * dv.aff.n, "affine assembly: element handle of another order");
    if (dv.cdof == 2) asm_gather_kernel<2, 2, COUNT, true><<<grid, 256, 0, st>>>(tv, dv, list, nlist, cnt, indptr, indices, data);
    else asm_gather_kernel<2, 1, COUNT, true><<<grid, 256, 0, st>>>(tv, dv, list, nlist, cnt, indptr, indices, data);
    PYN_CHECK_LAUNCH();
    return PYN_OK;
  }
#define PYN_ASM_CASE(R, C)                                                                                     \
  if (dv.rdof == R && dv.cdof == C) {                                                                          \
    asm_gather_kernel<R, C, COUNT><<<grid, 256, 0, st>>>(tv, dv, list, nlist, cnt, indptr, indices, data);     \
    PYN_CHECK_LAUNCH();                                                                                        \
    return PYN_OK;                                                                                             \
  }
  PYN_ASM_CASE(2, 2) PYN_ASM_CASE(2, 1) PYN_ASM_CASE(1, 2) PYN_ASM_CASE(3, 2) PYN_ASM_CASE(2, 3)
  PYN_ASM_CASE(1, 1) PYN_ASM_CASE(3, 3) PYN_ASM_CASE(1, 3) PYN_ASM_CASE(3, 1)
#undef PYN_ASM_CASE
  pyn::set_error("assembly: unsupported block shape %d x %d", dv.rdof, dv.cdof);
  return PYN_ERR_ARG;
}

static int launch_interior(const pyn_topo* t, const DescView& dv, const int64_t* indptr, int* indices, double* data,
                           cudaStream_t st) {
  const pyn::TopoView tv = pyn::view_of(t);
  const int grid = pyn::grid_for((t->nnodes + ASM_NB - 1) / ASM_NB * 256, 256, 8);
  if (dv.coef && !indices && data && dv.tile_desc) {   // structured mesh, values only: signature-grouped kernel
    PYN_REQUIRE(tv.nn == dv.aff.n * dv.aff.n, "affine assembly: element handle of another order");
    int64_t c0 = 0;
#define PYN_SIG_CASE(C, DG, cls, KERNEL, SHAPE)                                                                         \
    if (dv.cdof == C && dv.tile_class_end[cls] > c0) {                                                                  \
      const int64_t c1 = dv.tile_class_end[cls], g = (c1 - c0 + SIG_W - 1) / SIG_W;                                    \
      const size_t dyn = sizeof(double) * (size_t)SIG_W * SHAPE<C, DG>::per_warp;                                      \
      PYN_REQUIRE(g < 0x7fffffff, "affine assembly: too many signature tiles");                                        \
      PYN_CHECK_CUDA(cudaFuncSetAttribute(KERNEL<C, DG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));      \
      KERNEL<C, DG><<<(int)g, SIG_W * 32, dyn, st>>>(tv, dv, indptr, data, c0, c1,                                     \
                                                     dv.tile_slot + dv.tile_slot_base[cls]);                           \
      pyn::count_launch();                                                                                             \
    }
    // classes 0 / 2 / 4: valence 1 / 2 / <= 4, row-relative pieces; 1 / 3: valence 1 / 2 with line-aligned rows
    static const bool unaligned = getenv("PYN_SIG_UNALIGNED") != nullptr;   // A/B switch: row-relative pieces everywhere
    PYN_SIG_CASE(2, 1, 0, asm_sig_affine_kernel, SigShape) PYN_SIG_CASE(1, 1, 0, asm_sig_affine_kernel, SigShape)
    c0 = dv.tile_class_end[0];
    if (unaligned) {
      PYN_SIG_CASE(2, 1, 1, asm_sig_affine_kernel, SigShape) PYN_SIG_CASE(1, 1, 1, asm_sig_affine_kernel, SigShape)
    } else {
      PYN_SIG_CASE(2, 1, 1, asm_sigrow_affine_kernel, SigRowShape)
      PYN_SIG_CASE(1, 1, 1, asm_sigrow_affine_kernel, SigRowShape)
    }
    c0 = dv.tile_class_end[1];
    PYN_SIG_CASE(2, 2, 2, asm_sig_affine_kernel, SigShape) PYN_SIG_CASE(1, 2, 2, asm_sig_affine_kernel, SigShape)
    c0 = dv.tile_class_end[2];
    if (unaligned) {
      PYN_SIG_CASE(2, 2, 3, asm_sig_affine_kernel, SigShape) PYN_SIG_CASE(1, 2, 3, asm_sig_affine_kernel, SigShape)
    } else {
      PYN_SIG_CASE(2, 2, 3, asm_sigrow_affine_kernel, SigRowShape)
      PYN_SIG_CASE(1, 2, 3, asm_sigrow_affine_kernel, SigRowShape)
    }
    c0 = dv.tile_class_end[3];
    PYN_SIG_CASE(2, 4, 4, asm_sig_affine_kernel, SigShape) PYN_SIG_CASE(1, 4, 4, asm_sig_affine_kernel, SigShape)
#undef PYN_SIG_CASE
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
      pyn::set_error("signature-tile assembly: %s", cudaGetErrorString(e));
      return PYN_ERR_CUDA;
    }
    return PYN_OK;
  }
  if (dv.coef) {
    PYN_REQUIRE(tv.nn == dv.aff.n * dv.aff.n, "affine assembly: element handle of another order");
    // The product table is read through L1 (4 CTAs of 64 registers per SM: 6.1 ms for K at 707^2 x ngl 5); holding it
    // in shared memory instead costs a CTA per SM and measured 7.2 ms (PYN_AFF_TSM=1: A/B switch).
    const size_t tbytes = sizeof(double) * (size_t)tv.nn * tv.nn * (dv.cdof == 2 ? pyn::PYN_AFF_PK : pyn::PYN_AFF_PR);
    static const int tsm_env = getenv("PYN_AFF_TSM") ? atoi(getenv("PYN_AFF_TSM")) : 0;
    const bool tsm = tsm_env != 0 && tbytes <= 160 * 1024;
    const size_t dyn = tsm ? tbytes : 0;
#define PYN_AFF_LAUNCH(C, T)                                                                                         \
    {                                                                                                                  \
      PYN_CHECK_CUDA(cudaFuncSetAttribute(asm_interior_affine_kernel<C, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                          (int)dyn));                                                                  \
      asm_interior_affine_kernel<C, T><<<grid, 256, dyn, st>>>(tv, dv, indptr, indices, data);                         \
    }
    if (dv.cdof == 2) {
      if (tsm) PYN_AFF_LAUNCH(2, true) else PYN_AFF_LAUNCH(2, false)
    } else {
      if (tsm) PYN_AFF_LAUNCH(1, true) else PYN_AFF_LAUNCH(1, false)
    }
#undef PYN_AFF_LAUNCH
    PYN_CHECK_LAUNCH();
    return PYN_OK;
  }
  const bool vec = ((reinterpret_cast<uintptr_t>(dv.E) & 15) == 0) && ((dv.E_stride & 1) == 0);
#define PYN_ASM_CASE(R, C)                                                                        \
  if (dv.rdof == R && dv.cdof == C) {                                                             \
    if (C == 2 && vec)                                                                            \
      asm_interior_kernel<R, C, (C == 2)><<<grid, 256, 0, st>>>(tv, dv, indptr, indices, data);   \
    else                                                                                          \
      asm_interior_kernel<R, C, false><<<grid, 256, 0, st>>>(tv, dv, indptr, indices, data);      \
    PYN_CHECK_LAUNCH();                                                                           \
    return PYN_OK;                                                                                \
  }
  PYN_ASM_CASE(2, 2) PYN_ASM_CASE(2, 1) PYN_ASM_CASE(1, 2) PYN_ASM_CASE(3, 2) PYN_ASM_CASE(2, 3)
  PYN_ASM_CASE(1, 1) PYN_ASM_CASE(3, 3) PYN_ASM_CASE(1, 3) PYN_ASM_CASE(3, 1)
#undef PYN_ASM_CASE
  pyn::set_error("assembly: unsupported block shape %d x %d", dv.rdof, dv.cdof);
  return PYN_ERR_ARG;
}

static bool g_defer_join = false;    // pyn_asm_defer_join(1): a batch of fills joins the side stream once (pyn_asm_join)

// values-only assembly of the interior nodes through row templates (needs signatures and one element matrix)
static int launch_templates(const pyn_topo* t, const DescView& dv, const int64_t* indptr, double* data, cudaStream_t st) {
  const pyn::TopoView tv = pyn::view_of(t);
  pyn_topo* tm = const_cast<pyn_topo*>(t);   // scratch owned by the handle (one host thread per handle)
  if (!tm->sig_tmpl) PYN_CHECK_CUDA(cudaMalloc(&tm->sig_tmpl, sizeof(double) * (size_t)t->sig_slots * 9));
  double* T = tm->sig_tmpl;
  cudaStream_t tst = st;          // stream of the template kernel
  int ring = -1;
  static const bool no_overlap = getenv("PYN_ASM_NO_OVERLAP") != nullptr;
  if (g_defer_join && !no_overlap) {   // inside a batch: templates on the second side stream (topo.cuh)
    if (!tm->side2) {
      PYN_CHECK_CUDA(cudaStreamCreateWithFlags(&tm->side2, cudaStreamNonBlocking));
      PYN_CHECK_CUDA(cudaEventCreateWithFlags(&tm->ev_batch, cudaEventDisableTiming));
      for (int i = 0; i < 4; ++i) {
        PYN_CHECK_CUDA(cudaEventCreateWithFlags(&tm->ev_tmpl[i], cudaEventDisableTiming));
        PYN_CHECK_CUDA(cudaMalloc(&tm->sig_ring[i], sizeof(double) * (size_t)t->sig_slots * 9));
      }
    }
    if (!tm->batch_forked) {   // once per batch: the element matrices are ready at this point of the caller's stream
      PYN_CHECK_CUDA(cudaEventRecord(tm->ev_batch, st));
      PYN_CHECK_CUDA(cudaStreamWaitEvent(tm->side2, tm->ev_batch, 0));
      tm->batch_forked = 1;
      tm->ring_next = 0;
    }
    if (tm->ring_next < 4) {   // more than four template matrices in one batch: the rest stay on the caller's stream
      ring = tm->ring_next++;
      T = tm->sig_ring[ring];
      tst = tm->side2;
    }
  }
  const int tgrid = (t->nsig + 3) / 4;
  // One CTA per batch, NOT a persistent grid: CTAs are dispatched in batch order and start staggered, so the stores in
  // flight sweep the array like a plain fill.  Persistent CTAs stay in lock-step at addresses one batch stride apart and
  // lose ~20 % of the write bandwidth (tools/microbench/store_patterns*.cu: 6.2 vs 7.6 TB/s).  PYN_ASM_PERSISTENT=1: A/B.
  const int64_t nbatch_s = (t->nnodes + ASM_SB - 1) / ASM_SB;
  static const bool persistent_grid = getenv("PYN_ASM_PERSISTENT") != nullptr;
  const int sgrid = persistent_grid || nbatch_s > 0x7fffffff ? pyn::grid_for(nbatch_s * 256, 256, 8) : (int)nbatch_s;
  bool done = false;
  static const bool warp_only_set = [] {
    const int v = getenv("PYN_ASM_WARP_ONLY") ? 1 : 0;
    if (v) cudaMemcpyToSymbol(g_stream_warp_only, &v, sizeof(v));
    return true;
  }();
  (void)warp_only_set;
#define PYN_ASM_CASE(R, C)                                                                        \
  if (!done && dv.rdof == R && dv.cdof == C) {                                                    \
    asm_template_kernel<R, C><<<tgrid, 128, 0, tst>>>(tv, dv, T);                                 \
    pyn::count_launch();                                                                          \
    if (ring >= 0) {                                                                              \
      cudaEventRecord(tm->ev_tmpl[ring], tst);                                                    \
      cudaStreamWaitEvent(st, tm->ev_tmpl[ring], 0);                                              \
    }                                                                                             \
    asm_stream_kernel<R, C><<<sgrid, 256, 0, st>>>(tv, dv, indptr, T, data);                      \
    pyn::count_launch();                                                                          \
    done = true;                                                                                  \
  }
  PYN_ASM_CASE(2, 2) PYN_ASM_CASE(2, 1) PYN_ASM_CASE(1, 2) PYN_ASM_CASE(3, 2) PYN_ASM_CASE(2, 3)
  PYN_ASM_CASE(1, 1) PYN_ASM_CASE(3, 3) PYN_ASM_CASE(1, 3) PYN_ASM_CASE(3, 1)
#undef PYN_ASM_CASE
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    pyn::set_error("template assembly: %s", cudaGetErrorString(e));
    return PYN_ERR_CUDA;
  }
  if (!done) {
    pyn::set_error("assembly: unsupported block shape %d x %d", dv.rdof, dv.cdof);
    return PYN_ERR_ARG;
  }
  return PYN_OK;
}

// ---- signature tiles of a boundary map (lazy; see topo.cuh) ----------------------------------------------------------
namespace {
struct NotFlag {
  __host__ __device__ __forceinline__ uint8_t operator()(const uint8_t& f) const { return f ? 0 : 1; }
};
constexpr int SIG_CLASSES = 5;   // 2 * valence index + aligned; valence 4 (index 2) has no aligned variant
struct TileClasses {
  int64_t end[SIG_CLASSES];    // tiles of class c: [end[c-1], end[c])
  int64_t base[SIG_CLASSES];   // first int of the class in the slot table
};
__host__ __device__ inline int sig_valence_index(int cls) { return cls >> 1; }
// A tile takes the line-aligned kernel when its rows are long enough for the extra work to pay (measured: from 25 column
// nodes on) and hold an ODD number of column nodes: then every second row of a 2x2-block matrix starts 16 bytes into a
// sector and every second row of a 2x1-block matrix on an odd double (scalar stores).  Rows with an even number of
// column nodes lose with it (ngl 6, Rw: 1.42 -> 1.63 ms).
__host__ __device__ inline bool sig_wants_alignment(int deg, int L, int /*cdof*/) {
  return deg <= 2 && L >= 25 && (L & 1) != 0;
}
// group key of an un-flagged node: its row signature and where its first row starts inside a 128-byte line (in
// doubles: the rows of a signature have one length, so the second row's offset follows)
__global__ void sig_keys_kernel(pyn::TopoView t, int64_t n, const int* __restrict__ nodes,
                                const int64_t* __restrict__ indptr, int rdof, int cdof, int* __restrict__ keys,
                                int* __restrict__ counts) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = nodes[i];
    const int deg = (int)(t.n2e_ptr[j + 1] - t.n2e_ptr[j]), L = (int)(t.adj_ptr[j + 1] - t.adj_ptr[j]);
    const int cls = sig_wants_alignment(deg, L, cdof) ? (int)(indptr[(int64_t)j * rdof] & 15) : 0;
    const int s = t.sig_id[j] * 16 + cls;
    keys[i] = s;
    atomicAdd(&counts[s], 1);
  }
}
__global__ void tile_first_kernel(pyn::TopoView t, int64_t nt, const int* __restrict__ start,
                                  const int* __restrict__ sig_nodes, int* __restrict__ first, int* __restrict__ deg,
                                  int* __restrict__ len) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < nt; i += (int64_t)gridDim.x * blockDim.x) {
    const int j = sig_nodes[start[i]];
    first[i] = j;
    deg[i] = (int)(t.n2e_ptr[j + 1] - t.n2e_ptr[j]);
    len[i] = (int)(t.adj_ptr[j + 1] - t.adj_ptr[j]);
  }
}
__global__ void tile_build_kernel(pyn::TopoView t, int64_t nt, TileClasses tc, const int* __restrict__ tile_sig,
                                  const int* __restrict__ tile_start, const int* __restrict__ tile_cnt,
                                  const int* __restrict__ sig_nodes, const int64_t* __restrict__ indptr,
                                  int* __restrict__ desc, int* __restrict__ slot) {
  const int lane = threadIdx.x & 31;
  const int64_t w0 = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t i = w0; i < nt; i += nw) {
    int cls = 0;
    while (cls < SIG_CLASSES - 1 && i >= tc.end[cls]) ++cls;
    const int vi = sig_valence_index(cls);
    const int sl = 2 << vi, dmax = vi == 2 ? 4 : vi + 1;             // ints per slot, elements per node of the class
    const int64_t c0 = cls == 0 ? 0 : tc.end[cls - 1];
    const int start = tile_start[i], cnt = tile_cnt[i];
    const int j = sig_nodes[start + (lane < cnt ? lane : 0)];
    const int64_t b0 = t.n2e_ptr[j];
    const int dj = (int)(t.n2e_ptr[j + 1] - b0);
    int* sp = slot + tc.base[cls] + ((i - c0) * 32 + lane) * sl;
    sp[0] = j;
    for (int k = 0; k < sl - 1; ++k) sp[1 + k] = k < dmax ? t.n2e_elem[b0 + (k < dj ? k : 0)] : 0;
    if (lane == 0) {
      const int rep = t.sig_rep[tile_sig[i] >> 4];
      const int64_t rb0 = t.n2e_ptr[rep];
      const int deg = (int)(t.n2e_ptr[rep + 1] - rb0);
      unsigned lr = 0;
      for (int k = 0; k < deg && k < 4; ++k) lr |= (unsigned)(t.n2e_loc[rb0 + k] & 255) << (8 * k);
      const int64_t io = t.inv_ptr[rep];
      const int L = (int)(t.adj_ptr[rep + 1] - t.adj_ptr[rep]);
      int* o = desc + i * 4;
      o[0] = cnt | (deg << 8) | (L << 16);
      o[1] = (int)lr;
      o[2] = (int)(io & 0xffffffffll);
      const int tau0 = (int)(indptr[(int64_t)j * 2] & 15), tau1 = (int)(indptr[(int64_t)j * 2 + 1] & 15);
      o[3] = (int)((io >> 32) & 0xffffff) | (tau0 << 24) | (tau1 << 28);   // (inverse maps stay far below 2^56 bytes)
    }
  }
}
}  // namespace

static int ensure_sig_tiles(const pyn_topo* t, const pyn_bcmap* mc, const int64_t* indptr, int cdof, cudaStream_t st) {
  pyn_bcmap* m = const_cast<pyn_bcmap*>(mc);   // cache owned by the handle (one host thread per handle)
  if (m->tiles_state < 0 || (m->tiles_state == 1 && m->tiles_indptr == indptr)) return PYN_OK;
  if (m->tiles_state == 1) {   // the tiles follow the row starts of ONE matrix: another one came along with this map
    cudaFree(m->tile_desc); cudaFree(m->tile_slot);
    m->tile_desc = m->tile_slot = nullptr;
  }
  m->tiles_indptr = indptr;
  m->tiles_state = -1;
  // default on structured meshes; PYN_AFF_NO_SIG=1 keeps the flat kernel (A/B switch, and what the parity test uses to
  // cover both)
  static const bool off = getenv("PYN_AFF_NO_SIG") != nullptr;
  if (off || t->nsig <= 0 || t->max_valence > SIG_KMAX || t->nn > 255 || (int64_t)t->nn * SIG_KMAX >= 32768 ||
      t->nsig >= (1 << 26))
    return PYN_OK;   // (a row holds at most max_valence * nn column nodes: L fits the descriptor's 15 bits)
  const int64_t n = t->nnodes, nint = t->nnodes - m->nlist;
  if (nint <= 0 || nint >= 0x7fffffff || (int64_t)t->ne * t->nn >= 0x7fffffff) return PYN_OK;
  int *nodes = nullptr, *keys = nullptr, *keys2 = nullptr, *counts = nullptr;
  int64_t* d_n = nullptr;
  void* tmp = nullptr;
  size_t b1 = 0, b2 = 0;
  PYN_CHECK_CUDA(cudaMalloc(&nodes, sizeof(int) * nint));
  PYN_CHECK_CUDA(cudaMalloc(&m->sig_nodes, sizeof(int) * nint));
  PYN_CHECK_CUDA(cudaMalloc(&keys, sizeof(int) * nint));
  PYN_CHECK_CUDA(cudaMalloc(&keys2, sizeof(int) * nint));
  const int ngroups = t->nsig * 16;
  PYN_CHECK_CUDA(cudaMalloc(&counts, sizeof(int) * ngroups));
  PYN_CHECK_CUDA(cudaMalloc(&d_n, sizeof(int64_t)));
  PYN_CHECK_CUDA(cudaMemsetAsync(counts, 0, sizeof(int) * ngroups, st));
  cub::CountingInputIterator<int> ids(0);
  cub::TransformInputIterator<uint8_t, NotFlag, const uint8_t*> keep(m->flags, NotFlag());
  PYN_CHECK_CUDA(cub::DeviceSelect::Flagged(nullptr, b1, ids, keep, nodes, d_n, n, st));
  int end_bit = 1;
  while ((1 << end_bit) < ngroups) ++end_bit;
  PYN_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, b2, keys, keys2, nodes, m->sig_nodes, nint, 0, end_bit, st));
  PYN_CHECK_CUDA(cudaMalloc(&tmp, b1 > b2 ? b1 : b2));
  PYN_CHECK_CUDA(cub::DeviceSelect::Flagged(tmp, b1, ids, keep, nodes, d_n, n, st));      // ascending node ids
  sig_keys_kernel<<<pyn::grid_for(nint, 256), 256, 0, st>>>(pyn::view_of(t), nint, nodes, indptr, 2, cdof, keys, counts);
  PYN_CHECK_CUDA(cub::DeviceRadixSort::SortPairs(tmp, b2, keys, keys2, nodes, m->sig_nodes, nint, 0, end_bit, st));   // stable
  pyn::count_launch(3);
  std::vector<int> h_counts((size_t)ngroups);
  PYN_CHECK_CUDA(cudaMemcpyAsync(h_counts.data(), counts, sizeof(int) * ngroups, cudaMemcpyDeviceToHost, st));
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  cudaFree(nodes); cudaFree(keys); cudaFree(keys2); cudaFree(counts); cudaFree(d_n); cudaFree(tmp);
  std::vector<int> ts, tb, tc;
  int64_t off_ = 0;
  for (int s = 0; s < ngroups; ++s) {
    for (int c = 0; c < h_counts[s]; c += 32) {
      ts.push_back(s); tb.push_back((int)(off_ + c)); tc.push_back(h_counts[s] - c < 32 ? h_counts[s] - c : 32);
    }
    off_ += h_counts[s];
  }
  if (off_ != nint || ts.empty()) return PYN_OK;
  {  // order of execution: valence class (one launch each), then first node -- tiles that write neighbouring rows run
     // at about the same time, so the partial lines of a row meet in L2
    const size_t nt = ts.size();
    int *d_start = nullptr, *d_first = nullptr, *d_deg = nullptr, *d_len = nullptr;
    PYN_CHECK_CUDA(cudaMalloc(&d_start, sizeof(int) * nt));
    PYN_CHECK_CUDA(cudaMalloc(&d_first, sizeof(int) * nt));
    PYN_CHECK_CUDA(cudaMalloc(&d_deg, sizeof(int) * nt));
    PYN_CHECK_CUDA(cudaMalloc(&d_len, sizeof(int) * nt));
    PYN_CHECK_CUDA(cudaMemcpyAsync(d_start, tb.data(), sizeof(int) * nt, cudaMemcpyHostToDevice, st));
    tile_first_kernel<<<pyn::grid_for((int64_t)nt, 256), 256, 0, st>>>(pyn::view_of(t), (int64_t)nt, d_start, m->sig_nodes,
                                                                        d_first, d_deg, d_len);
    pyn::count_launch();
    std::vector<int> first(nt), dg(nt), ln(nt);
    PYN_CHECK_CUDA(cudaMemcpyAsync(first.data(), d_first, sizeof(int) * nt, cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaMemcpyAsync(dg.data(), d_deg, sizeof(int) * nt, cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaMemcpyAsync(ln.data(), d_len, sizeof(int) * nt, cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaStreamSynchronize(st));
    cudaFree(d_start); cudaFree(d_first); cudaFree(d_deg); cudaFree(d_len);
    std::vector<int64_t> key(nt);
    std::vector<int> order(nt);
    for (size_t i = 0; i < nt; ++i) {
      const int vi = dg[i] <= 1 ? 0 : (dg[i] == 2 ? 1 : 2);
      const int cls = 2 * vi + (sig_wants_alignment(dg[i], ln[i], cdof) ? 1 : 0);
      key[i] = ((int64_t)cls << 40) | (int64_t)first[i];
      order[i] = (int)i;
    }
    std::sort(order.begin(), order.end(), [&](int a, int b) { return key[a] < key[b]; });
    std::vector<int> ts2(nt), tb2(nt), tc2(nt);
    for (int c = 0; c < SIG_CLASSES; ++c) m->tile_class_end[c] = 0;
    for (size_t i = 0; i < nt; ++i) {
      const int o = order[i];
      ts2[i] = ts[o]; tb2[i] = tb[o]; tc2[i] = tc[o];
      m->tile_class_end[key[o] >> 40] = (int64_t)i + 1;
    }
    for (int c = 1; c < SIG_CLASSES; ++c)
      if (m->tile_class_end[c] < m->tile_class_end[c - 1]) m->tile_class_end[c] = m->tile_class_end[c - 1];
    ts.swap(ts2); tb.swap(tb2); tc.swap(tc2);
  }
  m->ntiles = (int64_t)ts.size();
  PYN_CHECK_CUDA(cudaMalloc(&m->tile_sig, sizeof(int) * ts.size()));
  PYN_CHECK_CUDA(cudaMalloc(&m->tile_start, sizeof(int) * ts.size()));
  PYN_CHECK_CUDA(cudaMalloc(&m->tile_cnt, sizeof(int) * ts.size()));
  PYN_CHECK_CUDA(cudaMemcpyAsync(m->tile_sig, ts.data(), sizeof(int) * ts.size(), cudaMemcpyHostToDevice, st));
  PYN_CHECK_CUDA(cudaMemcpyAsync(m->tile_start, tb.data(), sizeof(int) * ts.size(), cudaMemcpyHostToDevice, st));
  PYN_CHECK_CUDA(cudaMemcpyAsync(m->tile_cnt, tc.data(), sizeof(int) * ts.size(), cudaMemcpyHostToDevice, st));
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  TileClasses tcl;
  int64_t slot_ints = 0;
  for (int c = 0; c < SIG_CLASSES; ++c) {
    tcl.end[c] = m->tile_class_end[c];
    tcl.base[c] = m->tile_slot_base[c] = slot_ints;
    slot_ints += (tcl.end[c] - (c ? tcl.end[c - 1] : 0)) * 32 * (2 << sig_valence_index(c));
  }
  PYN_CHECK_CUDA(cudaMalloc(&m->tile_desc, sizeof(int) * 4 * ts.size()));
  PYN_CHECK_CUDA(cudaMalloc(&m->tile_slot, sizeof(int) * (slot_ints > 0 ? slot_ints : 1)));
  tile_build_kernel<<<pyn::grid_for((int64_t)ts.size() * 32, 256), 256, 0, st>>>(
      pyn::view_of(t), (int64_t)ts.size(), tcl, m->tile_sig, m->tile_start, m->tile_cnt, m->sig_nodes, indptr,
      m->tile_desc, m->tile_slot);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  cudaFree(m->sig_nodes); cudaFree(m->tile_sig); cudaFree(m->tile_start); cudaFree(m->tile_cnt);
  m->sig_nodes = m->tile_sig = m->tile_start = m->tile_cnt = nullptr;
  m->tiles_state = 1;
  return PYN_OK;
}

// the interior fast path applies when a map is given and class-0 rows carry no forced diagonal
static bool g_no_templates = false;  // pyn_asm_set_templates(0): force the generic gather kernel (tests, comparisons)
static inline bool use_map(const pyn_asm_desc* desc) { return desc->bcmap != nullptr && !(desc->diag_classes & 1u); }

extern "C" {

int pyn_bcmap_create(const pyn_topo* t, const int8_t* d_rclass, int rdof, const int8_t* d_cclass, int cdof,
                     void* stream, pyn_bcmap** out) {
  PYN_REQUIRE(t && out && rdof >= 0 && rdof <= 3 && cdof >= 0 && cdof <= 3, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  pyn_bcmap* m = new pyn_bcmap();
  memset(m, 0, sizeof(*m));
  m->nnodes = t->nnodes;
  m->st = st;
  PYN_CHECK_CUDA(cudaMallocAsync(&m->flags, (size_t)t->nnodes, st));
  const pyn::TopoView tv = pyn::view_of(t);
  bcmap_own_kernel<<<pyn::grid_for(t->nnodes, 256), 256, 0, st>>>(tv, d_rclass, rdof, m->flags);
  PYN_CHECK_LAUNCH();
  if (d_cclass && cdof > 0) {
    bcmap_scatter_kernel<<<pyn::grid_for(t->nnodes, 256), 256, 0, st>>>(tv, d_cclass, cdof, m->flags);
    PYN_CHECK_LAUNCH();
  }
  // ascending list of the flagged nodes: stream-ordered temporaries, one host read-back (the list length)
  int* sel = nullptr;
  int64_t* d_n = nullptr;
  void* tmp = nullptr;
  size_t bytes = 0;
  cub::CountingInputIterator<int> ids(0);
  PYN_CHECK_CUDA(cub::DeviceSelect::Flagged(tmp, bytes, ids, m->flags, sel, d_n, t->nnodes, st));
  PYN_CHECK_CUDA(cudaMallocAsync(&sel, sizeof(int) * t->nnodes, st));
  PYN_CHECK_CUDA(cudaMallocAsync(&d_n, sizeof(int64_t), st));
  PYN_CHECK_CUDA(cudaMallocAsync(&tmp, bytes, st));
  PYN_CHECK_CUDA(cub::DeviceSelect::Flagged(tmp, bytes, ids, m->flags, sel, d_n, t->nnodes, st));
  pyn::count_launch(2);
  {
    const int rc = read_device_i64(d_n, &m->nlist, st);
    if (rc) return rc;
  }
  PYN_CHECK_CUDA(cudaMallocAsync(&m->list, sizeof(int) * (m->nlist > 0 ? m->nlist : 1), st));
  PYN_CHECK_CUDA(cudaMemcpyAsync(m->list, sel, sizeof(int) * m->nlist, cudaMemcpyDeviceToDevice, st));
  PYN_CHECK_CUDA(cudaFreeAsync(tmp, st));
  PYN_CHECK_CUDA(cudaFreeAsync(sel, st));
  PYN_CHECK_CUDA(cudaFreeAsync(d_n, st));
  *out = m;
  return PYN_OK;
}

int pyn_bcmap_destroy(pyn_bcmap* m) {
  if (!m) return PYN_OK;
  cudaFreeAsync(m->flags, m->st);   // stream-ordered: no device-wide synchronisation, the pool keeps the memory
  cudaFreeAsync(m->list, m->st);
  cudaFree(m->sig_nodes); cudaFree(m->tile_sig); cudaFree(m->tile_start); cudaFree(m->tile_cnt);
  cudaFree(m->tile_desc); cudaFree(m->tile_slot);
  delete m;
  return PYN_OK;
}

int pyn_bcmap_info(const pyn_bcmap* m, int64_t* n) {
  PYN_REQUIRE(m, "bcmap is NULL");
  if (n) *n = m->nlist;
  return PYN_OK;
}

int pyn_asm_count(const pyn_topo* t, const pyn_asm_desc* desc, int64_t* d_indptr, int64_t* h_nnz, void* stream) {
  PYN_REQUIRE(t && d_indptr, "bad arguments");
  DescView dv;
  int rc = make_desc(desc, &dv);
  if (rc) return rc;
  cudaStream_t st = pyn::as_stream(stream);
  const int64_t nrows = t->nnodes * dv.rdof;
  int* cnt = nullptr;
  PYN_CHECK_CUDA(cudaMallocAsync(&cnt, sizeof(int) * (nrows + 1), st));
  if (use_map(desc)) {
    PYN_REQUIRE(desc->bcmap->nnodes == t->nnodes, "bcmap belongs to another topology");
    asm_interior_count_kernel<<<pyn::grid_for(t->nnodes, 256), 256, 0, st>>>(pyn::view_of(t), dv.flags, dv.rdof, dv.cdof,
                                                                             (int)(dv.pair_mask & 1u), cnt);
    PYN_CHECK_LAUNCH();
    rc = launch_asm<true>(t, dv, desc->bcmap->list, desc->bcmap->nlist, cnt, nullptr, nullptr, nullptr, st);
  } else {
    rc = launch_asm<true>(t, dv, nullptr, 0, cnt, nullptr, nullptr, nullptr, st);
  }
  if (rc == PYN_OK) rc = pyn::scan_counts(cnt, nrows, d_indptr, h_nnz, st);
  PYN_CHECK_CUDA(cudaFreeAsync(cnt, st));
  return rc;
}

// boundary-list kernel of a values-only fill on the topology's side stream, ordered after the caller's stream so far
static int fork_boundary(const pyn_topo* t, const DescView& dv, const int* list, int64_t nlist, const int64_t* indptr,
                         double* data, cudaStream_t st) {
  pyn_topo* tm = const_cast<pyn_topo*>(t);   // one host thread per handle
  if (!tm->side) {
    PYN_CHECK_CUDA(cudaStreamCreateWithFlags(&tm->side, cudaStreamNonBlocking));
    PYN_CHECK_CUDA(cudaEventCreateWithFlags(&tm->ev_fork, cudaEventDisableTiming));
    PYN_CHECK_CUDA(cudaEventCreateWithFlags(&tm->ev_join, cudaEventDisableTiming));
  }
  PYN_CHECK_CUDA(cudaEventRecord(tm->ev_fork, st));
  PYN_CHECK_CUDA(cudaStreamWaitEvent(tm->side, tm->ev_fork, 0));
  const int rc = launch_asm<false>(t, dv, list, nlist, nullptr, indptr, nullptr, data, tm->side);
  if (rc) return rc;
  PYN_CHECK_CUDA(cudaEventRecord(tm->ev_join, tm->side));
  tm->pending = 1;
  return PYN_OK;
}

int pyn_asm_join(const pyn_topo* t, void* stream) {
  PYN_REQUIRE(t, "topo is NULL");
  pyn_topo* tm = const_cast<pyn_topo*>(t);
  tm->batch_forked = 0;   // the next batch forks its template stream again (its element matrices are new)
  if (tm->pending) {
    PYN_CHECK_CUDA(cudaStreamWaitEvent(pyn::as_stream(stream), tm->ev_join, 0));
    tm->pending = 0;
  }
  return PYN_OK;
}

int pyn_asm_defer_join(int enabled) {
  g_defer_join = enabled != 0;
  return PYN_OK;
}

int pyn_asm_fill(const pyn_topo* t, const pyn_asm_desc* desc, const int64_t* d_indptr, int32_t* d_indices,
                 double* d_data, void* stream) {
  PYN_REQUIRE(t && d_indptr, "bad arguments");
  DescView dv;
  int rc = make_desc(desc, &dv);
  if (rc) return rc;
  PYN_REQUIRE(d_data == nullptr || dv.E != nullptr || dv.coef != nullptr, "element matrices missing");
  cudaStream_t st = pyn::as_stream(stream);
  if (use_map(desc)) {
    PYN_REQUIRE(desc->bcmap->nnodes == t->nnodes, "bcmap belongs to another topology");
    if ((dv.pair_mask & 1u) && (d_data || d_indices)) {
      // values on a known pattern, one element matrix for all cells, repetitive mesh: stream the row templates
      const bool tmpl = d_data && !d_indices && dv.E_stride == 0 && t->nsig > 0 && !g_no_templates;
      // values-only re-assembly: the boundary list goes to the side stream first, so that it overlaps the interior
      // rows written on the caller's stream (disjoint rows; both read the same element matrix)
      const bool overlap = d_data && !d_indices && desc->bcmap->nlist > 0 && !getenv("PYN_ASM_NO_OVERLAP");
      if (overlap) {
        rc = fork_boundary(t, dv, desc->bcmap->list, desc->bcmap->nlist, d_indptr, d_data, st);
        if (rc) return rc;
      }
      if (dv.coef && d_data && !d_indices && dv.rdof == 2) {   // structured mesh: the signature-grouped fused kernel
        rc = ensure_sig_tiles(t, desc->bcmap, d_indptr, dv.cdof, st);
        if (rc) return rc;
        if (desc->bcmap->tiles_state == 1) {
          dv.tile_desc = desc->bcmap->tile_desc; dv.tile_slot = desc->bcmap->tile_slot; dv.ntiles = desc->bcmap->ntiles;
          for (int c = 0; c < SIG_CLASSES; ++c) {
            dv.tile_class_end[c] = desc->bcmap->tile_class_end[c];
            dv.tile_slot_base[c] = desc->bcmap->tile_slot_base[c];
          }
        }
      }
      rc = tmpl ? launch_templates(t, dv, d_indptr, d_data, st) : launch_interior(t, dv, d_indptr, d_indices, d_data, st);
      if (rc) return rc;
      if (overlap) return g_defer_join ? PYN_OK : pyn_asm_join(t, stream);
    } else if (d_data && !d_indices && desc->bcmap->nlist > 0 && !getenv("PYN_ASM_NO_OVERLAP")) {
      // boundary-only matrix inside a batch: side stream as well, joined with the batch
      rc = fork_boundary(t, dv, desc->bcmap->list, desc->bcmap->nlist, d_indptr, d_data, st);
      if (rc) return rc;
      return g_defer_join ? PYN_OK : pyn_asm_join(t, stream);
    }
    return launch_asm<false>(t, dv, desc->bcmap->list, desc->bcmap->nlist, nullptr, d_indptr, d_indices, d_data, st);
  }
  return launch_asm<false>(t, dv, nullptr, 0, nullptr, d_indptr, d_indices, d_data, st);
}

int pyn_asm_set_templates(int enabled) {
  g_no_templates = !enabled;
  return PYN_OK;
}

int pyn_topo_signatures(const pyn_topo* t, int* nsig, int64_t* slots) {
  PYN_REQUIRE(t, "topo is NULL");
  if (nsig) *nsig = t->nsig;
  if (slots) *slots = t->sig_slots;
  return PYN_OK;
}

int pyn_tmat_create(const pyn_topo* t, const pyn_asm_desc* desc, const int64_t* d_indptr, const int32_t* d_indices,
                    const double* d_data, void* stream, pyn_tmat** out) {
  PYN_REQUIRE(t && desc && d_indptr && d_indices && d_data && out, "bad arguments");
  DescView dv;
  int rc = make_desc(desc, &dv);
  if (rc) return rc;
  PYN_REQUIRE(dv.E && dv.E_stride == 0, "element operator: needs ONE element matrix for all cells (E_stride == 0)");
  PYN_REQUIRE(use_map(desc) && (dv.pair_mask & 1u), "element operator: needs a boundary map and the (free,free) pair");
  PYN_REQUIRE(!(dv.diag_classes & 1u), "element operator: a diagonal rule on free dofs is not supported");
  PYN_REQUIRE(desc->bcmap->nnodes == t->nnodes, "bcmap belongs to another topology");
  PYN_REQUIRE(pyn_tmat_supported(t->nn, dv.rdof, dv.cdof), "element operator: element matrix too large for shared memory");
  pyn_tmat* m = new pyn_tmat();
  memset(m, 0, sizeof(*m));
  m->topo = t; m->rdof = dv.rdof; m->cdof = dv.cdof;
  m->flags = desc->bcmap->flags; m->list = desc->bcmap->list; m->nlist = desc->bcmap->nlist;
  m->rowscale = dv.rowscale; m->ptr = d_indptr; m->ind = d_indices; m->val = d_data;
  rc = pyn::tmat_init(m, dv.E, dv.sign, pyn::as_stream(stream));
  if (rc) { pyn_tmat_destroy(m); return rc; }
  *out = m;
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

int pyn_csr_diag_offsets(int64_t nrows, const int64_t* p, const int32_t* ind, int32_t* off, void* stream) {
  PYN_REQUIRE(nrows > 0 && p && ind && off, "bad arguments");
  csr_diag_offsets_kernel<<<pyn::grid_for(nrows, 256), 256, 0, pyn::as_stream(stream)>>>(nrows, p, ind, off);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_csr_diagonal_at(int64_t nrows, const int64_t* p, const int32_t* off, const double* val, double* diag,
                        void* stream) {
  PYN_REQUIRE(nrows > 0 && p && off && val && diag, "bad arguments");
  csr_diag_at_kernel<<<pyn::grid_for(nrows, 256), 256, 0, pyn::as_stream(stream)>>>(nrows, p, off, val, diag);
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
