// Mesh numbering on device + node->element incidence + node adjacency (pattern skeleton shared by all matrices).
#include <cub/cub.cuh>

#include "common.cuh"
#include "topo.cuh"

namespace {

// ---- serial DMPlexCreateBoxMesh numbering (see oracle/kle_oracle.py BoxMesh) -----------
__global__ void box_cell_nodes_kernel(int nx, int ny, int n, int64_t cell0, int64_t ncells, int* __restrict__ out) {
  const int nn = n * n, m = n - 2;
  const int64_t tot = ncells * nn;
  const int64_t ncell_all = (int64_t)nx * ny;
  const int64_t off_vert = ncell_all * m * m;
  const int64_t off_edge = off_vert + (int64_t)(nx + 1) * (ny + 1);
  const int64_t nxedge = (int64_t)nx * (ny + 1);
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < tot; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t cell = cell0 + idx / nn;
    const int k = (int)(idx % nn);
    const int64_t fx = cell % nx, fy = cell / nx;
    int64_t node;
    if (k < 4) {  // vertices in closure order
      const int64_t vx = fx + ((k == 1 || k == 2) ? 1 : 0), vy = fy + ((k >= 2) ? 1 : 0);
      node = off_vert + vy * (nx + 1) + vx;
    } else if (k < 4 + 4 * m) {
      const int ed = (k - 4) / m, t = (k - 4) % m;
      if (ed == 0) node = off_edge + (fy * nx + fx) * m + t;                              // bottom, +x
      else if (ed == 1) node = off_edge + (nxedge + (fx + 1) * ny + fy) * m + t;          // right, +y
      else if (ed == 2) node = off_edge + ((fy + 1) * nx + fx) * m + (m - 1 - t);         // top, reversed
      else node = off_edge + (nxedge + fx * ny + fy) * m + (m - 1 - t);                   // left, reversed
    } else {
      node = cell * m * m + (k - 4 - 4 * m);
    }
    out[idx] = (int)node;
  }
}

__global__ void box_cell_corners_kernel(int nx, int ny, double x0, double y0, double hx, double hy, int64_t cell0,
                                        int64_t ncells, double* __restrict__ out) {
  for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < ncells; c += (int64_t)gridDim.x * blockDim.x) {
    const int64_t cell = cell0 + c;
    const double fx = (double)(cell % nx), fy = (double)(cell / nx);
    const double xa = x0 + hx * fx, xb = x0 + hx * (fx + 1.0), ya = y0 + hy * fy, yb = y0 + hy * (fy + 1.0);
    double* o = out + c * 8;
    o[0] = xa; o[1] = ya; o[2] = xb; o[3] = ya; o[4] = xb; o[5] = yb; o[6] = xa; o[7] = yb;
  }
}

struct XiPerm {
  double xi[PYN_MAX_NGL];
  int perm[PYN_MAX_NGL * PYN_MAX_NGL];
};

__global__ void box_node_coords_kernel(int nx, int ny, int n, double x0, double y0, double hx, double hy,
                                       const __grid_constant__ XiPerm tp, int64_t fy_off,
                                       double* __restrict__ xy) {
  // one thread per (cell, local node); later cells overwrite shared nodes with the same value up to rounding,
  // so to stay deterministic only the LAST cell containing a node (reference: INSERT in cell order) writes it.
  const int nn = n * n, m = n - 2;
  const int64_t ncell = (int64_t)nx * ny, tot = ncell * nn;
  const int64_t off_vert = ncell * m * m;
  const int64_t off_edge = off_vert + (int64_t)(nx + 1) * (ny + 1);
  const int64_t nxedge = (int64_t)nx * (ny + 1);
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < tot; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t cell = idx / nn;
    const int k = (int)(idx % nn);
    const int64_t fx = cell % nx, fy = cell / nx;
    int64_t node;
    bool last = true;  // is this the highest-numbered cell touching the node?
    if (k < 4) {
      const int64_t vx = fx + ((k == 1 || k == 2) ? 1 : 0), vy = fy + ((k >= 2) ? 1 : 0);
      node = off_vert + vy * (nx + 1) + vx;
      last = (fx == (vx < nx ? vx : nx - 1)) && (fy == (vy < ny ? vy : ny - 1));
    } else if (k < 4 + 4 * m) {
      const int ed = (k - 4) / m, t = (k - 4) % m;
      if (ed == 0) {         // bottom edge: also the top edge of the cell below, which comes earlier
        node = off_edge + (fy * nx + fx) * m + t;
      } else if (ed == 1) {  // right edge: also the left edge of the next cell in the row
        node = off_edge + (nxedge + (fx + 1) * ny + fy) * m + t;
        last = (fx == nx - 1);
      } else if (ed == 2) {  // top edge: also the bottom edge of the cell above
        node = off_edge + ((fy + 1) * nx + fx) * m + (m - 1 - t);
        last = (fy == ny - 1);
      } else {               // left edge: also the right edge of the previous cell
        node = off_edge + (nxedge + fx * ny + fy) * m + (m - 1 - t);
      }
    } else {
      node = cell * m * m + (k - 4 - 4 * m);
    }
    if (!last) continue;
    const int p = tp.perm[k];
    const double r = tp.xi[p / n], s = tp.xi[p % n];
    // bilinear map with corner k at (+1,+1),(-1,+1),(-1,-1),(+1,-1) and corners (xa,ya),(xb,ya),(xb,yb),(xa,yb)
    const double xa = x0 + hx * (double)fx, xb = x0 + hx * (double)(fx + 1);
    const double ya = y0 + hy * (double)(fy + fy_off), yb = y0 + hy * (double)(fy + fy_off + 1);
    const double rp = (r + 1.0) * 0.5, rm = (1.0 - r) * 0.5, sp = (s + 1.0) * 0.5, sm_ = (1.0 - s) * 0.5;
    const double N0 = rp * sp, N1 = rm * sp, N2 = rm * sm_, N3 = rp * sm_;
    xy[2 * node] = N0 * xa + N1 * xb + N2 * xb + N3 * xa;
    xy[2 * node + 1] = N0 * ya + N1 * ya + N2 * yb + N3 * yb;
  }
}

// ---- node -> element incidence ---------------------------------------------------------------------------
__global__ void count_valence_kernel(const int* __restrict__ e2n, int64_t tot, int* __restrict__ cnt) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x)
    atomicAdd(&cnt[e2n[i]], 1);
}
__global__ void fill_incidence_kernel(const int* __restrict__ e2n, int64_t tot, int nn, const int64_t* __restrict__ ptr,
                                      int* __restrict__ cursor, int* __restrict__ elems, short* __restrict__ locs) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x) {
    const int node = e2n[i];
    const int slot = atomicAdd(&cursor[node], 1);
    elems[ptr[node] + slot] = (int)(i / nn);
    locs[ptr[node] + slot] = (short)(i % nn);
  }
}
// ascending element order per node (insertion sort; valence is tiny) -> deterministic summation order
__global__ void sort_incidence_kernel(int64_t nnodes, const int64_t* __restrict__ ptr, int* __restrict__ elems,
                                      short* __restrict__ locs) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = ptr[r], e = ptr[r + 1];
    for (int64_t i = b + 1; i < e; ++i) {
      const int ke = elems[i];
      const short kl = locs[i];
      int64_t j = i - 1;
      while (j >= b && elems[j] > ke) {
        elems[j + 1] = elems[j];
        locs[j + 1] = locs[j];
        --j;
      }
      elems[j + 1] = ke;
      locs[j + 1] = kl;
    }
  }
}

// ---- node adjacency: one warp per node, bitonic sort of the candidate list in shared memory ----------------
template <bool FILL>
__global__ void __launch_bounds__(128)
adjacency_kernel(pyn::TopoView t, int cap, int* __restrict__ rowlen, const int64_t* __restrict__ rowptr,
                 int* __restrict__ cols) {
  extern __shared__ int smi[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  int* buf = smi + w * cap;
  for (int64_t r = blockIdx.x * (int64_t)wpb + w; r < t.nnodes; r += (int64_t)gridDim.x * wpb) {
    const int64_t b = t.n2e_ptr[r];
    const int deg = (int)(t.n2e_ptr[r + 1] - b);
    const int cnt = deg * t.nn;
    int P = 32;  // sort only the power of two that holds this node's candidates
    while (P < cnt) P <<= 1;
    for (int i = lane; i < P; i += 32) {
      int v = 0x7fffffff;
      if (i < cnt) v = t.e2n[(int64_t)t.n2e_elem[b + i / t.nn] * t.nn + (i % t.nn)];
      buf[i] = v;
    }
    __syncwarp();
    for (int k = 2; k <= P; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int i = lane; i < P; i += 32) {
          const int ixj = i ^ j;
          if (ixj > i) {
            const int a = buf[i], c = buf[ixj];
            const bool up = ((i & k) == 0);
            if ((a > c) == up) { buf[i] = c; buf[ixj] = a; }
          }
        }
        __syncwarp();
      }
    }
    // unique count / compaction
    int base = 0;
    const int64_t out0 = FILL ? rowptr[r] : 0;
    for (int i0 = 0; i0 < cnt; i0 += 32) {
      const int i = i0 + lane;
      const bool keep = (i < cnt) && (i == 0 || buf[i] != buf[i - 1]);
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (FILL && keep) cols[out0 + base + __popc(m & ((1u << lane) - 1))] = buf[i];
      base += __popc(m);
    }
    if (!FILL && lane == 0) rowlen[r] = base;
    __syncwarp();
  }
}

// per node: valence * adjacency row length (size of its slice of the inverse position map)
__global__ void inv_count_kernel(pyn::TopoView t, int* __restrict__ cnt) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < t.nnodes; r += (int64_t)gridDim.x * blockDim.x)
    cnt[r] = (int)(t.n2e_ptr[r + 1] - t.n2e_ptr[r]) * (int)(t.adj_ptr[r + 1] - t.adj_ptr[r]);
}

// one warp per node: inv[k][p] = local index in element k of adjacency column p (binary search of every element node
// in the sorted adjacency row held in shared memory)
__global__ void __launch_bounds__(128)
inv_fill_kernel(pyn::TopoView t, int cap, uint8_t* __restrict__ inv) {
  extern __shared__ int smi[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  int* row = smi + w * cap;
  for (int64_t r = blockIdx.x * (int64_t)wpb + w; r < t.nnodes; r += (int64_t)gridDim.x * wpb) {
    const int64_t a0 = t.adj_ptr[r], b0 = t.n2e_ptr[r], o = t.inv_ptr[r];
    const int L = (int)(t.adj_ptr[r + 1] - a0), deg = (int)(t.n2e_ptr[r + 1] - b0);
    for (int i = lane; i < L; i += 32) row[i] = t.adj_cols[a0 + i];
    for (int i = lane; i < deg * L; i += 32) inv[o + i] = 0xFF;
    __syncwarp();
    for (int i = lane; i < deg * t.nn; i += 32) {
      const int k = i / t.nn, l = i % t.nn;
      const int c = t.e2n[(int64_t)t.n2e_elem[b0 + k] * t.nn + l];
      int lo = 0, hi = L;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (row[mid] < c) lo = mid + 1; else hi = mid;
      }
      inv[o + (int64_t)k * L + lo] = (uint8_t)l;
    }
    __syncwarp();
  }
}

__global__ void int_to_i64_kernel(const int* __restrict__ in, int64_t n, int64_t* __restrict__ out) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    out[i] = in[i];
}

}  // namespace

namespace pyn {
// exclusive scan of int counts into int64 row pointers (ptr[n] = total); returns total on the host
int scan_counts(const int* d_cnt, int64_t n, int64_t* d_ptr, int64_t* h_total, cudaStream_t st) {
  int64_t* tmp64 = nullptr;
  PYN_CHECK_CUDA(cudaMallocAsync(&tmp64, sizeof(int64_t) * (n + 1), st));
  int_to_i64_kernel<<<grid_for(n, 256), 256, 0, st>>>(d_cnt, n, tmp64);
  PYN_CHECK_LAUNCH();
  PYN_CHECK_CUDA(cudaMemsetAsync(tmp64 + n, 0, sizeof(int64_t), st));
  void* d_temp = nullptr;
  size_t bytes = 0;
  PYN_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(d_temp, bytes, tmp64, d_ptr, n + 1, st));
  PYN_CHECK_CUDA(cudaMallocAsync(&d_temp, bytes, st));
  PYN_CHECK_CUDA(cub::DeviceScan::ExclusiveSum(d_temp, bytes, tmp64, d_ptr, n + 1, st));
  count_launch(2);
  if (h_total) {
    PYN_CHECK_CUDA(cudaMemcpyAsync(h_total, d_ptr + n, sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  }
  PYN_CHECK_CUDA(cudaFreeAsync(d_temp, st));
  PYN_CHECK_CUDA(cudaFreeAsync(tmp64, st));
  return PYN_OK;
}

// ---- row signatures ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t sig_mix(uint64_t h, uint64_t v) {
  h ^= v;
  h *= 0x100000001b3ULL;   // FNV-1a step on 64-bit words
  h ^= h >> 29;
  return h;
}
__global__ void sig_hash_kernel(pyn::TopoView t, uint64_t* __restrict__ key, int* __restrict__ node) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < t.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b0 = t.n2e_ptr[r];
    const int deg = (int)(t.n2e_ptr[r + 1] - b0), L = (int)(t.adj_ptr[r + 1] - t.adj_ptr[r]);
    uint64_t h = sig_mix(0xcbf29ce484222325ULL, ((uint64_t)deg << 32) | (uint32_t)L);
    for (int k = 0; k < deg; ++k) h = sig_mix(h, (uint64_t)(uint16_t)t.n2e_loc[b0 + k] + 0x9e37ULL * (k + 1));
    const uint8_t* iv = t.inv + t.inv_ptr[r];
    const int len = deg * L;
    uint64_t w = 0;
    for (int i = 0; i < len; ++i) {
      w = (w << 8) | iv[i];
      if ((i & 7) == 7) { h = sig_mix(h, w); w = 0; }
    }
    h = sig_mix(h, w + 0x51ULL);
    key[r] = h;
    node[r] = (int)r;
  }
}
__global__ void sig_heads_kernel(int64_t n, const uint64_t* __restrict__ key, int* __restrict__ head) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    head[i] = (i == 0 || key[i] != key[i - 1]) ? 1 : 0;
}
// gid = inclusive scan of heads - 1 ; the sort is stable, so the head of a group is its lowest-numbered node
__global__ void sig_assign_kernel(int64_t n, const int* __restrict__ head, const int64_t* __restrict__ scan,
                                  const int* __restrict__ node, int* __restrict__ sig_id, int* __restrict__ rep) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(scan[i + 1] - 1);
    sig_id[node[i]] = g;
    if (head[i]) rep[g] = node[i];
  }
}
// exact comparison with the representative (a hash collision disables the fast path for this mesh)
__global__ void sig_verify_kernel(pyn::TopoView t, const int* __restrict__ sig_id, const int* __restrict__ rep,
                                  int* __restrict__ bad) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < t.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t q = rep[sig_id[r]];
    if (q == r) continue;
    const int64_t b0 = t.n2e_ptr[r], c0 = t.n2e_ptr[q];
    const int deg = (int)(t.n2e_ptr[r + 1] - b0), L = (int)(t.adj_ptr[r + 1] - t.adj_ptr[r]);
    bool same = deg == (int)(t.n2e_ptr[q + 1] - c0) && L == (int)(t.adj_ptr[q + 1] - t.adj_ptr[q]);
    if (same) {
      for (int k = 0; k < deg; ++k) same &= t.n2e_loc[b0 + k] == t.n2e_loc[c0 + k];
      const uint8_t *a = t.inv + t.inv_ptr[r], *b = t.inv + t.inv_ptr[q];
      for (int i = 0; i < deg * L && same; ++i) same = a[i] == b[i];
    }
    if (!same) atomicAdd(bad, 1);
  }
}
__global__ void sig_len_kernel(int nsig, pyn::TopoView t, const int* __restrict__ rep, int* __restrict__ len) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nsig; i += gridDim.x * blockDim.x)
    len[i] = (int)(t.adj_ptr[rep[i] + 1] - t.adj_ptr[rep[i]]);
}

// Builds sig_id / sig_rep / sig_ptr; leaves nsig = 0 when the mesh has too many distinct rows to make templates pay.
int build_signatures(pyn_topo* t, cudaStream_t st) {
  const int64_t n = t->nnodes;
  t->nsig = 0;
  if (n >= 2147483647LL) return PYN_OK;
  uint64_t *key = nullptr, *key2 = nullptr;
  int *node = nullptr, *node2 = nullptr, *head = nullptr, *bad = nullptr, *len = nullptr;
  int64_t* scan = nullptr;
  void* tmp = nullptr;
  size_t bytes = 0;
  int rc = PYN_OK;
  auto fail = [&](const char* what, cudaError_t e) {
    pyn::set_error("build_signatures: %s: %s", what, cudaGetErrorString(e));
    rc = PYN_ERR_CUDA;
  };
  cudaError_t e;
  const pyn::TopoView v = pyn::view_of(t);
  do {
    if ((e = cudaMalloc(&key, 8 * n)) || (e = cudaMalloc(&key2, 8 * n)) || (e = cudaMalloc(&node, 4 * n)) ||
        (e = cudaMalloc(&node2, 4 * n)) || (e = cudaMalloc(&head, 4 * (n + 1))) || (e = cudaMalloc(&scan, 8 * (n + 1))) ||
        (e = cudaMalloc(&bad, 4)) || (e = cudaMalloc(&t->sig_id, 4 * n))) { fail("malloc", e); break; }
    sig_hash_kernel<<<pyn::grid_for(n, 128, 32), 128, 0, st>>>(v, key, node);
    pyn::count_launch();
    cub::DeviceRadixSort::SortPairs(tmp, bytes, key, key2, node, node2, n, 0, 64, st);
    if ((e = cudaMalloc(&tmp, bytes))) { fail("malloc sort", e); break; }
    cub::DeviceRadixSort::SortPairs(tmp, bytes, key, key2, node, node2, n, 0, 64, st);
    pyn::count_launch();
    sig_heads_kernel<<<pyn::grid_for(n, 256), 256, 0, st>>>(n, key2, head);
    pyn::count_launch();
    int64_t nsig64 = 0;
    rc = pyn::scan_counts(head, n, scan, &nsig64, st);
    if (rc) break;
    if (nsig64 > 4096 || nsig64 * 8 > n) break;      // not a repetitive mesh: keep the generic kernels
    const int nsig = (int)nsig64;
    if ((e = cudaMalloc(&t->sig_rep, 4 * nsig)) || (e = cudaMalloc(&t->sig_ptr, 8 * (nsig + 1))) ||
        (e = cudaMalloc(&len, 4 * (nsig + 1)))) { fail("malloc rep", e); break; }
    sig_assign_kernel<<<pyn::grid_for(n, 256), 256, 0, st>>>(n, head, scan, node2, t->sig_id, t->sig_rep);
    pyn::count_launch();
    cudaMemsetAsync(bad, 0, 4, st);
    sig_verify_kernel<<<pyn::grid_for(n, 128, 32), 128, 0, st>>>(v, t->sig_id, t->sig_rep, bad);
    pyn::count_launch();
    int h_bad = 0;
    if ((e = cudaMemcpyAsync(&h_bad, bad, 4, cudaMemcpyDeviceToHost, st)) || (e = cudaStreamSynchronize(st))) {
      fail("verify", e);
      break;
    }
    if (h_bad) break;                                  // hash collision: no templates for this mesh
    sig_len_kernel<<<pyn::grid_for(nsig, 256), 256, 0, st>>>(nsig, v, t->sig_rep, len);
    pyn::count_launch();
    rc = pyn::scan_counts(len, nsig, t->sig_ptr, &t->sig_slots, st);
    if (rc) break;
    t->nsig = nsig;
  } while (0);
  cudaFree(key); cudaFree(key2); cudaFree(node); cudaFree(node2); cudaFree(head); cudaFree(scan); cudaFree(bad);
  cudaFree(len); cudaFree(tmp);
  return rc;
}
}  // namespace pyn

extern "C" {

int pyn_box_cell_nodes(int nx, int ny, int ngl, int64_t cell0, int64_t ncells, int32_t* d_out, void* stream) {
  PYN_REQUIRE(nx > 0 && ny > 0 && ngl >= 2 && d_out, "bad arguments");
  const int m = ngl - 2;
  const double nn_tot = (double)nx * ny * m * m + (double)(nx + 1) * (ny + 1) +
                        ((double)nx * (ny + 1) + (double)(nx + 1) * ny) * m;
  if (nn_tot >= 2147483647.0) {
    pyn::set_error("pyn_box_cell_nodes: %g nodes do not fit int32 node ids", nn_tot);
    return PYN_ERR_LIMIT;
  }
  if (ncells == 0) return PYN_OK;
  box_cell_nodes_kernel<<<pyn::grid_for(ncells * ngl * ngl, 256), 256, 0, pyn::as_stream(stream)>>>(
      nx, ny, ngl, cell0, ncells, d_out);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_box_cell_corners(int nx, int ny, double x0, double y0, double x1, double y1, int64_t cell0, int64_t ncells,
                         double* d_coords, void* stream) {
  PYN_REQUIRE(nx > 0 && ny > 0 && d_coords, "bad arguments");
  if (ncells == 0) return PYN_OK;
  box_cell_corners_kernel<<<pyn::grid_for(ncells, 256), 256, 0, pyn::as_stream(stream)>>>(
      nx, ny, x0, y0, (x1 - x0) / nx, (y1 - y0) / ny, cell0, ncells, d_coords);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_box_node_coords(int nx, int ny, int ngl, double x0, double y0, double x1, double y1, const double* h_xi,
                        const int32_t* h_perm, double* d_xy, void* stream) {
  PYN_REQUIRE(nx > 0 && ny > 0 && ngl >= 2 && ngl <= PYN_MAX_NGL && h_xi && h_perm && d_xy, "bad arguments");
  XiPerm tp;
  for (int i = 0; i < ngl; ++i) tp.xi[i] = h_xi[i];
  for (int i = 0; i < ngl * ngl; ++i) tp.perm[i] = h_perm[i];
  box_node_coords_kernel<<<pyn::grid_for((int64_t)nx * ny * ngl * ngl, 256), 256, 0, pyn::as_stream(stream)>>>(
      nx, ny, ngl, x0, y0, (x1 - x0) / nx, (y1 - y0) / ny, tp, 0, d_xy);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_box_node_coords_strip(int nx, int ny_global, int64_t fy0, int ny_local, int ngl, double x0, double y0, double x1,
                              double y1, const double* h_xi, const int32_t* h_perm, double* d_xy, void* stream) {
  PYN_REQUIRE(nx > 0 && ny_global > 0 && ny_local > 0 && fy0 >= 0 && fy0 + ny_local <= ny_global && ngl >= 2 &&
                  ngl <= PYN_MAX_NGL && h_xi && h_perm && d_xy, "bad arguments");
  XiPerm tp;
  for (int i = 0; i < ngl; ++i) tp.xi[i] = h_xi[i];
  for (int i = 0; i < ngl * ngl; ++i) tp.perm[i] = h_perm[i];
  box_node_coords_kernel<<<pyn::grid_for((int64_t)nx * ny_local * ngl * ngl, 256), 256, 0, pyn::as_stream(stream)>>>(
      nx, ny_local, ngl, x0, y0, (x1 - x0) / nx, (y1 - y0) / ny_global, tp, fy0, d_xy);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

int pyn_topo_create(const int32_t* d_e2n, int64_t ne, int nn, int64_t nnodes, void* stream, pyn_topo** out) {
  PYN_REQUIRE(d_e2n && ne > 0 && nn > 0 && nnodes > 0 && out, "bad arguments");
  cudaStream_t st = pyn::as_stream(stream);
  pyn_topo* t = new pyn_topo();
  memset(t, 0, sizeof(*t));
  t->ne = ne; t->nn = nn; t->nnodes = nnodes; t->e2n = d_e2n;
  const int64_t tot = ne * nn;
  int* cnt = nullptr;
  int rc = PYN_OK;
#define TOPO_TRY(x) do { rc = (x); if (rc != PYN_OK) { pyn_topo_destroy(t); return rc; } } while (0)
  auto cuda_try = [&](cudaError_t e, const char* what) -> int {
    if (e != cudaSuccess) { pyn::set_error("pyn_topo_create: %s: %s", what, cudaGetErrorString(e)); return PYN_ERR_CUDA; }
    return PYN_OK;
  };
  TOPO_TRY(cuda_try(cudaMalloc(&cnt, sizeof(int) * (nnodes + 1)), "malloc cnt"));
  TOPO_TRY(cuda_try(cudaMemsetAsync(cnt, 0, sizeof(int) * (nnodes + 1), st), "memset"));
  count_valence_kernel<<<pyn::grid_for(tot, 256), 256, 0, st>>>(d_e2n, tot, cnt);
  pyn::count_launch();
  TOPO_TRY(cuda_try(cudaMalloc(&t->n2e_ptr, sizeof(int64_t) * (nnodes + 1)), "malloc n2e_ptr"));
  int64_t total = 0;
  TOPO_TRY(pyn::scan_counts(cnt, nnodes, t->n2e_ptr, &total, st));
  // max valence
  {
    void* d_temp = nullptr; size_t bytes = 0; int* d_max = nullptr;
    TOPO_TRY(cuda_try(cudaMalloc(&d_max, sizeof(int)), "malloc"));
    cub::DeviceReduce::Max(d_temp, bytes, cnt, d_max, nnodes, st);
    TOPO_TRY(cuda_try(cudaMalloc(&d_temp, bytes), "malloc"));
    cub::DeviceReduce::Max(d_temp, bytes, cnt, d_max, nnodes, st);
    pyn::count_launch();
    TOPO_TRY(cuda_try(cudaMemcpyAsync(&t->max_valence, d_max, sizeof(int), cudaMemcpyDeviceToHost, st), "memcpy"));
    TOPO_TRY(cuda_try(cudaStreamSynchronize(st), "sync"));
    cudaFree(d_temp); cudaFree(d_max);
  }
  TOPO_TRY(cuda_try(cudaMalloc(&t->n2e_elem, sizeof(int) * total), "malloc n2e_elem"));
  TOPO_TRY(cuda_try(cudaMalloc(&t->n2e_loc, sizeof(short) * total), "malloc n2e_loc"));
  TOPO_TRY(cuda_try(cudaMemsetAsync(cnt, 0, sizeof(int) * (nnodes + 1), st), "memset"));
  fill_incidence_kernel<<<pyn::grid_for(tot, 256), 256, 0, st>>>(d_e2n, tot, nn, t->n2e_ptr, cnt, t->n2e_elem, t->n2e_loc);
  pyn::count_launch();
  sort_incidence_kernel<<<pyn::grid_for(nnodes, 256), 256, 0, st>>>(nnodes, t->n2e_ptr, t->n2e_elem, t->n2e_loc);
  pyn::count_launch();
  // adjacency
  int cap = 32;
  while (cap < t->max_valence * nn) cap <<= 1;
  if ((size_t)cap * 4 * sizeof(int) > 200 * 1024) {
    pyn::set_error("pyn_topo_create: valence %d x %d nodes/element exceeds the adjacency sort capacity", t->max_valence, nn);
    pyn_topo_destroy(t); cudaFree(cnt);
    return PYN_ERR_LIMIT;
  }
  pyn::TopoView v = pyn::view_of(t);
  const size_t smem = (size_t)cap * 4 * sizeof(int);
  cudaFuncSetAttribute(adjacency_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(adjacency_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int grid = pyn::grid_for(nnodes, 4, 16);
  adjacency_kernel<false><<<grid, 128, smem, st>>>(v, cap, cnt, nullptr, nullptr);
  pyn::count_launch();
  TOPO_TRY(cuda_try(cudaMalloc(&t->adj_ptr, sizeof(int64_t) * (nnodes + 1)), "malloc adj_ptr"));
  TOPO_TRY(pyn::scan_counts(cnt, nnodes, t->adj_ptr, &t->adj_nnz, st));
  {
    void* d_temp = nullptr; size_t bytes = 0; int* d_max = nullptr;
    TOPO_TRY(cuda_try(cudaMalloc(&d_max, sizeof(int)), "malloc"));
    cub::DeviceReduce::Max(d_temp, bytes, cnt, d_max, nnodes, st);
    TOPO_TRY(cuda_try(cudaMalloc(&d_temp, bytes), "malloc"));
    cub::DeviceReduce::Max(d_temp, bytes, cnt, d_max, nnodes, st);
    pyn::count_launch();
    TOPO_TRY(cuda_try(cudaMemcpyAsync(&t->max_row, d_max, sizeof(int), cudaMemcpyDeviceToHost, st), "memcpy"));
    TOPO_TRY(cuda_try(cudaStreamSynchronize(st), "sync"));
    cudaFree(d_temp); cudaFree(d_max);
  }
  TOPO_TRY(cuda_try(cudaMalloc(&t->adj_cols, sizeof(int) * t->adj_nnz), "malloc adj_cols"));
  adjacency_kernel<true><<<grid, 128, smem, st>>>(v, cap, nullptr, t->adj_ptr, t->adj_cols);
  pyn::count_launch();
  // inverse position map for the gather assembly
  if (nn > 254 || t->max_valence > 32) {
    pyn::set_error("pyn_topo_create: %d nodes/element (max 254) or valence %d (max 32) outside the compiled limits", nn,
                   t->max_valence);
    pyn_topo_destroy(t); cudaFree(cnt);
    return PYN_ERR_LIMIT;
  }
  v = pyn::view_of(t);
  inv_count_kernel<<<pyn::grid_for(nnodes, 256), 256, 0, st>>>(v, cnt);
  pyn::count_launch();
  TOPO_TRY(cuda_try(cudaMalloc(&t->inv_ptr, sizeof(int64_t) * (nnodes + 1)), "malloc inv_ptr"));
  TOPO_TRY(pyn::scan_counts(cnt, nnodes, t->inv_ptr, &t->inv_len, st));
  TOPO_TRY(cuda_try(cudaMalloc(&t->inv, (size_t)t->inv_len + 32), "malloc inv"));
  v = pyn::view_of(t);
  {
    int rcap = 32;
    while (rcap < t->max_row) rcap <<= 1;
    const size_t smem2 = (size_t)rcap * 4 * sizeof(int);
    cudaFuncSetAttribute(inv_fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
    inv_fill_kernel<<<grid, 128, smem2, st>>>(v, rcap, t->inv);
    pyn::count_launch();
  }
  TOPO_TRY(cuda_try(cudaStreamSynchronize(st), "sync"));
  TOPO_TRY(cuda_try(cudaGetLastError(), "kernels"));
  cudaFree(cnt);
  TOPO_TRY(pyn::build_signatures(t, st));
#undef TOPO_TRY
  *out = t;
  return PYN_OK;
}

int pyn_topo_destroy(pyn_topo* t) {
  if (!t) return PYN_OK;
  cudaFree(t->n2e_ptr); cudaFree(t->n2e_elem); cudaFree(t->n2e_loc); cudaFree(t->adj_ptr); cudaFree(t->adj_cols); cudaFree(t->inv_ptr); cudaFree(t->inv);
  cudaFree(t->sig_id); cudaFree(t->sig_rep); cudaFree(t->sig_ptr); cudaFree(t->sig_tmpl);
  if (t->side) { cudaStreamSynchronize(t->side); cudaStreamDestroy(t->side); cudaEventDestroy(t->ev_fork); cudaEventDestroy(t->ev_join); }
  if (t->side2) {
    cudaStreamSynchronize(t->side2);
    cudaStreamDestroy(t->side2);
    cudaEventDestroy(t->ev_batch);
    for (int i = 0; i < 4; ++i) { cudaEventDestroy(t->ev_tmpl[i]); cudaFree(t->sig_ring[i]); }
  }
  delete t;
  return PYN_OK;
}

int pyn_topo_info(const pyn_topo* t, int64_t* adj_nnz, int* max_valence, int* max_row) {
  PYN_REQUIRE(t, "topo is NULL");
  if (adj_nnz) *adj_nnz = t->adj_nnz;
  if (max_valence) *max_valence = t->max_valence;
  if (max_row) *max_row = t->max_row;
  return PYN_OK;
}

int pyn_topo_adjacency(const pyn_topo* t, const int64_t** d_rowptr, const int32_t** d_cols) {
  PYN_REQUIRE(t, "topo is NULL");
  if (d_rowptr) *d_rowptr = t->adj_ptr;
  if (d_cols) *d_cols = t->adj_cols;
  return PYN_OK;
}
}
