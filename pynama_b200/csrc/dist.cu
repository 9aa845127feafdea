// Multi-GPU plumbing of the path: one process per GPU, NCCL over NVLink/NVSwitch.
//
// In the reference these roles are played by PETSc/MPI: the VecScatter inside MPIAIJ MatMult (-> halo
// exchange of interface-node dofs) and MPI_Allreduce inside VecDot/VecNorm (-> all-reduce of the Krylov scalars).
// NCCL is bound at run time (dlopen of libnccl.so.2, the copy PyTorch already loaded when there is one), so the
// library itself only needs libcudart and single-GPU users never touch NCCL.
#include <dlfcn.h>

#include <vector>

#include "dist.cuh"

namespace {

// ---- minimal NCCL ABI (nccl.h: stable since 2.x) ----------------------------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclFloat64 = 8 };   // ncclDataType_t: ncclDouble
enum { ncclSum = 0 };

struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  int (*GetVersion)(int*) = nullptr;
} g_nccl;

int load_nccl() {
  if (g_nccl.lib) return PYN_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* h = nullptr;
  for (const char* n : names) {
    h = dlopen(n, RTLD_NOW | RTLD_NOLOAD);   // prefer the copy already in the process (PyTorch's)
    if (h) break;
  }
  if (!h)
    for (const char* n : names) {
      h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (h) break;
    }
  if (!h) {
    pyn::set_error("NCCL not found: dlopen(libnccl.so.2) failed: %s", dlerror());
    return PYN_ERR_NCCL;
  }
#define SYM(field, name)                                                       \
  do {                                                                         \
    *(void**)(&g_nccl.field) = dlsym(h, name);                                 \
    if (!g_nccl.field) {                                                       \
      pyn::set_error("NCCL symbol %s missing", name);                          \
      return PYN_ERR_NCCL;                                                     \
    }                                                                          \
  } while (0)
  SYM(GetUniqueId, "ncclGetUniqueId");
  SYM(CommInitRank, "ncclCommInitRank");
  SYM(CommDestroy, "ncclCommDestroy");
  SYM(AllReduce, "ncclAllReduce");
  SYM(Send, "ncclSend");
  SYM(Recv, "ncclRecv");
  SYM(GroupStart, "ncclGroupStart");
  SYM(GroupEnd, "ncclGroupEnd");
  SYM(GetErrorString, "ncclGetErrorString");
  SYM(GetVersion, "ncclGetVersion");
#undef SYM
  g_nccl.lib = h;
  return PYN_OK;
}

#define PYN_CHECK_NCCL(expr)                                                                      \
  do {                                                                                            \
    int _r = (expr);                                                                              \
    if (_r != ncclSuccess) {                                                                      \
      pyn::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, g_nccl.GetErrorString(_r));    \
      return PYN_ERR_NCCL;                                                                        \
    }                                                                                             \
  } while (0)

// gather the listed nodes' dofs into a contiguous buffer / scatter them back
__global__ void halo_pack_kernel(int64_t n, int ncomp, const int64_t* __restrict__ idx, const double* __restrict__ v,
                                 double* __restrict__ buf) {
  const int64_t tot = n * ncomp;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x)
    buf[i] = v[idx[i / ncomp] * ncomp + (i % ncomp)];
}
__global__ void halo_unpack_kernel(int64_t n, int ncomp, const int64_t* __restrict__ idx, const double* __restrict__ buf,
                                   double* __restrict__ v) {
  const int64_t tot = n * ncomp;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x)
    v[idx[i / ncomp] * ncomp + (i % ncomp)] = buf[i];
}

// ---- peer-memory path -------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

struct PeerView {
  int rank, world;
  double* local;
  double* peer[PYN_PEER_MAX];
  int64_t cap;
  unsigned long long* ctr;
};
PeerView view_of(const pyn_peer* p) {
  PeerView v;
  v.rank = p->rank; v.world = p->world; v.local = p->local; v.cap = p->cap; v.ctr = p->ctr;
  for (int i = 0; i < PYN_PEER_MAX; ++i) v.peer[i] = p->peer[i];
  return v;
}
__host__ __device__ inline int64_t ar_doubles(int world) { return 2ll * world * PYN_PEER_SLOT; }
__host__ __device__ inline int64_t mbox_doubles(int64_t cap) { return 8 + cap; }
// mailbox (side, parity) of a region: side 0 = message from the rank below, 1 = from the rank above
__device__ __forceinline__ double* mbox(double* region, int world, int64_t cap, int side, int par) {
  return region + ar_doubles(world) + (int64_t)(side * 2 + par) * mbox_doubles(cap);
}

// One warp.  vals = local sums (from per-CTA partials, or the in-place buffer), pushed to every rank's slot `rank`,
// then summed over the slots in rank order.
__global__ void peer_allreduce_kernel(PeerView P, int na, const double* __restrict__ pa, int nb,
                                      const double* __restrict__ pb, int n, double* __restrict__ io) {
  const int lane = threadIdx.x;
  double vals[PYN_PEER_SLOT - 1];
#pragma unroll
  for (int i = 0; i < PYN_PEER_SLOT - 1; ++i) vals[i] = 0.0;
  if (pa) {
    vals[0] = pyn::reduce_partials(pa, na);
    if (pb) vals[1] = pyn::reduce_partials(pb, nb);
  } else {
#pragma unroll
    for (int i = 0; i < PYN_PEER_SLOT - 1; ++i)
      if (i < n) vals[i] = io[i];
  }
  const unsigned long long e = P.ctr[0] + 1ull;
  const int par = (int)(e & 1ull);
  if (lane < P.world) {
    double* slot = P.peer[lane] + ((int64_t)par * P.world + P.rank) * PYN_PEER_SLOT;
#pragma unroll
    for (int i = 0; i < PYN_PEER_SLOT - 1; ++i)
      if (i < n) slot[i] = vals[i];
    __threadfence_system();
    st_release_sys(reinterpret_cast<unsigned long long*>(slot + PYN_PEER_SLOT - 1), e);
    const double* mine = P.local + ((int64_t)par * P.world + lane) * PYN_PEER_SLOT;
    while (ld_acquire_sys(reinterpret_cast<const unsigned long long*>(mine + PYN_PEER_SLOT - 1)) < e) {
    }
  }
  __syncwarp();
  if (lane == 0) {
    __threadfence_system();
    for (int i = 0; i < n; ++i) {
      double s = 0.0;
      for (int r = 0; r < P.world; ++r)
        s += *reinterpret_cast<volatile const double*>(P.local + ((int64_t)par * P.world + r) * PYN_PEER_SLOT + i);
      io[i] = s;
    }
    P.ctr[0] = e;
  }
}

struct HaloP2P {
  int nnbr;
  int side_at_peer[PYN_HALO_MAX_NBR];   // which mailbox of the neighbour receives my message
  int side_local[PYN_HALO_MAX_NBR];     // which of my mailboxes receives the neighbour's
  int nbr[PYN_HALO_MAX_NBR];
  int64_t send_cnt[PYN_HALO_MAX_NBR], recv_cnt[PYN_HALO_MAX_NBR];
  const int64_t* send_idx[PYN_HALO_MAX_NBR];
  const int64_t* recv_idx[PYN_HALO_MAX_NBR];
};

// interface values straight into the neighbours' mailboxes (parity of the next epoch)
__global__ void halo_put_kernel(PeerView P, HaloP2P H, int ncomp, const double* __restrict__ v) {
  const unsigned long long e = P.ctr[1] + 1ull;
  const int par = (int)(e & 1ull);
  for (int k = 0; k < H.nnbr; ++k) {
    double* dst = mbox(P.peer[H.nbr[k]], P.world, P.cap, H.side_at_peer[k], par) + 8;
    const int64_t tot = H.send_cnt[k] * ncomp;
    const int64_t* idx = H.send_idx[k];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x)
      dst[i] = v[idx[i / ncomp] * ncomp + (i % ncomp)];
  }
  __threadfence_system();
  if (blockIdx.x == 0 && threadIdx.x == 0) P.ctr[2] = e;
}
// release the epoch flag at the neighbours, wait for theirs, scatter the received values into the ghost entries.
// The grid never exceeds the number of SMs, so every block is resident while block 0 publishes the flags.
__global__ void halo_wait_kernel(PeerView P, HaloP2P H, int ncomp, double* __restrict__ v) {
  const unsigned long long e = P.ctr[2];
  const int par = (int)(e & 1ull);
  if (blockIdx.x == 0 && threadIdx.x < H.nnbr) {
    const int k = threadIdx.x;
    __threadfence_system();
    st_release_sys(reinterpret_cast<unsigned long long*>(mbox(P.peer[H.nbr[k]], P.world, P.cap, H.side_at_peer[k], par)), e);
  }
  if (threadIdx.x == 0) {
    for (int k = 0; k < H.nnbr; ++k) {
      const unsigned long long* f =
          reinterpret_cast<const unsigned long long*>(mbox(P.local, P.world, P.cap, H.side_local[k], par));
      while (ld_acquire_sys(f) < e) {
      }
    }
    __threadfence_system();
  }
  __syncthreads();
  for (int k = 0; k < H.nnbr; ++k) {
    const double* src = mbox(P.local, P.world, P.cap, H.side_local[k], par) + 8;
    const int64_t tot = H.recv_cnt[k] * ncomp;
    const int64_t* idx = H.recv_idx[k];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < tot; i += (int64_t)gridDim.x * blockDim.x)
      v[idx[i / ncomp] * ncomp + (i % ncomp)] = *reinterpret_cast<volatile const double*>(src + i);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) P.ctr[1] = e;
}

__global__ void reduce2_kernel(int na, const double* __restrict__ pa, int nb, const double* __restrict__ pb,
                               double* __restrict__ out) {
  const double a = pyn::reduce_partials(pa, na);
  const double b = pb ? pyn::reduce_partials(pb, nb) : 0.0;
  if (threadIdx.x == 0) {
    out[0] = a;
    if (pb) out[1] = b;
  }
}

}  // namespace

namespace pyn {

int comm_allreduce_sum(const pyn_comm* c, double* d_buf, int n, cudaStream_t st) {
  if (!c || c->world == 1) return PYN_OK;
  if (c->peer && n < PYN_PEER_SLOT) {
    peer_allreduce_kernel<<<1, 32, 0, st>>>(view_of(c->peer), 0, nullptr, 0, nullptr, n, d_buf);
    PYN_CHECK_LAUNCH();
    return PYN_OK;
  }
  PYN_CHECK_NCCL(g_nccl.AllReduce(d_buf, d_buf, (size_t)n, ncclFloat64, ncclSum, (ncclComm_t)c->nccl, st));
  return PYN_OK;
}

int comm_reduce_allreduce(const pyn_comm* c, int na, const double* pa, int nb, const double* pb, double* d_out,
                          cudaStream_t st) {
  const int n = pb ? 2 : 1;
  if (c && c->world > 1 && c->peer) {
    peer_allreduce_kernel<<<1, 32, 0, st>>>(view_of(c->peer), na, pa, nb, pb, n, d_out);
    PYN_CHECK_LAUNCH();
    return PYN_OK;
  }
  reduce2_kernel<<<1, 32, 0, st>>>(na, pa, nb, pb, d_out);
  PYN_CHECK_LAUNCH();
  return comm_allreduce_sum(c, d_out, n, st);
}

static bool halo_p2p_ok(const pyn_halo* h, int ncomp) {
  const pyn_peer* p = h->comm->peer;
  if (!p || !p->opened) return false;
  for (int k = 0; k < h->nnbr; ++k) {
    if (h->nbr[k] != p->rank - 1 && h->nbr[k] != p->rank + 1) return false;   // strips: neighbours are rank +- 1
    if (h->send_cnt[k] * ncomp > p->cap || h->recv_cnt[k] * ncomp > p->cap) return false;
  }
  return true;
}

int halo_exchange(const pyn_halo* h, double* d_vec, int ncomp, cudaStream_t st) {
  if (!h || h->nnbr == 0) return PYN_OK;
  PYN_REQUIRE(ncomp >= 1 && ncomp <= PYN_HALO_MAX_COMP, "halo: ncomp out of range");
  if (halo_p2p_ok(h, ncomp)) {
    const pyn_peer* p = h->comm->peer;
    HaloP2P H;
    memset(&H, 0, sizeof(H));
    H.nnbr = h->nnbr;
    int64_t most = 1;
    for (int k = 0; k < h->nnbr; ++k) {
      const bool above = h->nbr[k] == p->rank + 1;
      H.nbr[k] = h->nbr[k];
      H.side_at_peer[k] = above ? 0 : 1;   // the rank above files me under "from below"
      H.side_local[k] = above ? 1 : 0;
      H.send_cnt[k] = h->send_cnt[k]; H.recv_cnt[k] = h->recv_cnt[k];
      H.send_idx[k] = h->send_idx[k]; H.recv_idx[k] = h->recv_idx[k];
      most = max(most, max(h->send_cnt[k], h->recv_cnt[k]) * ncomp);
    }
    int grid = (int)((most + 255) / 256);
    if (grid > pyn::sm_count()) grid = pyn::sm_count();
    const PeerView P = view_of(p);
    halo_put_kernel<<<grid, 256, 0, st>>>(P, H, ncomp, d_vec);
    PYN_CHECK_LAUNCH();
    halo_wait_kernel<<<grid, 256, 0, st>>>(P, H, ncomp, d_vec);
    PYN_CHECK_LAUNCH();
    return PYN_OK;
  }
  for (int k = 0; k < h->nnbr; ++k) {
    if (h->send_cnt[k] > 0) {
      halo_pack_kernel<<<pyn::grid_for(h->send_cnt[k] * ncomp, 256), 256, 0, st>>>(h->send_cnt[k], ncomp, h->send_idx[k],
                                                                                  d_vec, h->send_buf[k]);
      pyn::count_launch();
    }
  }
  PYN_CHECK_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < h->nnbr; ++k) {
    if (h->send_cnt[k] > 0)
      PYN_CHECK_NCCL(g_nccl.Send(h->send_buf[k], (size_t)(h->send_cnt[k] * ncomp), ncclFloat64, h->nbr[k],
                                 (ncclComm_t)h->comm->nccl, st));
    if (h->recv_cnt[k] > 0)
      PYN_CHECK_NCCL(g_nccl.Recv(h->recv_buf[k], (size_t)(h->recv_cnt[k] * ncomp), ncclFloat64, h->nbr[k],
                                 (ncclComm_t)h->comm->nccl, st));
  }
  PYN_CHECK_NCCL(g_nccl.GroupEnd());
  for (int k = 0; k < h->nnbr; ++k) {
    if (h->recv_cnt[k] > 0) {
      halo_unpack_kernel<<<pyn::grid_for(h->recv_cnt[k] * ncomp, 256), 256, 0, st>>>(h->recv_cnt[k], ncomp, h->recv_idx[k],
                                                                                    h->recv_buf[k], d_vec);
      pyn::count_launch();
    }
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    pyn::set_error("halo_exchange: %s", cudaGetErrorString(e));
    return PYN_ERR_CUDA;
  }
  return PYN_OK;
}

}  // namespace pyn

extern "C" {

int pyn_comm_unique_id(char* id128) {
  PYN_REQUIRE(id128, "id buffer is NULL");
  int rc = load_nccl();
  if (rc) return rc;
  ncclUniqueId id;
  PYN_CHECK_NCCL(g_nccl.GetUniqueId(&id));
  memcpy(id128, id.internal, 128);
  return PYN_OK;
}

int pyn_comm_create(const char* id128, int rank, int world, pyn_comm** out) {
  PYN_REQUIRE(id128 && out && world >= 1 && rank >= 0 && rank < world, "bad arguments");
  pyn_comm* c = new pyn_comm();
  c->rank = rank; c->world = world; c->nccl = nullptr; c->peer = nullptr;
  if (world > 1) {
    int rc = load_nccl();
    if (rc) { delete c; return rc; }
    ncclUniqueId id;
    memcpy(id.internal, id128, 128);
    ncclComm_t comm;
    int r = g_nccl.CommInitRank(&comm, world, id, rank);
    if (r != ncclSuccess) {
      pyn::set_error("ncclCommInitRank -> %s", g_nccl.GetErrorString(r));
      delete c;
      return PYN_ERR_NCCL;
    }
    c->nccl = comm;
  }
  *out = c;
  return PYN_OK;
}

int pyn_comm_destroy(pyn_comm* c) {
  if (!c) return PYN_OK;
  if (c->nccl) g_nccl.CommDestroy((ncclComm_t)c->nccl);
  delete c;
  return PYN_OK;
}

int pyn_comm_info(const pyn_comm* c, int* rank, int* world, int* nccl_version) {
  PYN_REQUIRE(c, "comm is NULL");
  if (rank) *rank = c->rank;
  if (world) *world = c->world;
  if (nccl_version) {
    *nccl_version = 0;
    if (g_nccl.lib) g_nccl.GetVersion(nccl_version);
  }
  return PYN_OK;
}

int pyn_comm_allreduce_sum(const pyn_comm* c, double* d_buf, int n, void* stream) {
  PYN_REQUIRE(c && d_buf && n > 0, "bad arguments");
  return pyn::comm_allreduce_sum(c, d_buf, n, pyn::as_stream(stream));
}

/* ---- peer memory (CUDA IPC over NVLink) ---- */
int pyn_peer_create(int rank, int world, int64_t cap_doubles, pyn_peer** out) {
  PYN_REQUIRE(out && world >= 2 && world <= PYN_PEER_MAX && rank >= 0 && rank < world && cap_doubles >= 0, "bad arguments");
  pyn_peer* p = new pyn_peer();
  memset(p, 0, sizeof(*p));
  p->rank = rank; p->world = world; p->cap = (cap_doubles + 1) & ~(int64_t)1;
  p->bytes = sizeof(double) * (size_t)(ar_doubles(world) + 4 * mbox_doubles(p->cap));
  PYN_CHECK_CUDA(cudaMalloc(&p->local, p->bytes));
  PYN_CHECK_CUDA(cudaMemset(p->local, 0, p->bytes));
  PYN_CHECK_CUDA(cudaMalloc(&p->ctr, 4 * sizeof(unsigned long long)));
  PYN_CHECK_CUDA(cudaMemset(p->ctr, 0, 4 * sizeof(unsigned long long)));
  PYN_CHECK_CUDA(cudaDeviceSynchronize());
  p->peer[rank] = p->local;
  *out = p;
  return PYN_OK;
}

int pyn_peer_handle(const pyn_peer* p, char* handle64) {
  PYN_REQUIRE(p && handle64, "bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  PYN_CHECK_CUDA(cudaIpcGetMemHandle(&h, p->local));
  memcpy(handle64, &h, 64);
  return PYN_OK;
}

/* handles: world x 64 bytes, rank-major (gathered by the caller's rendezvous layer) */
int pyn_peer_open(pyn_peer* p, const char* handles) {
  PYN_REQUIRE(p && handles, "bad arguments");
  for (int r = 0; r < p->world; ++r) {
    if (r == p->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + 64 * r, 64);
    void* ptr = nullptr;
    PYN_CHECK_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    p->peer[r] = static_cast<double*>(ptr);
  }
  p->opened = 1;
  return PYN_OK;
}

int pyn_peer_destroy(pyn_peer* p) {
  if (!p) return PYN_OK;
  for (int r = 0; r < p->world; ++r)
    if (r != p->rank && p->peer[r]) cudaIpcCloseMemHandle(p->peer[r]);
  cudaFree(p->local);
  cudaFree(p->ctr);
  delete p;
  return PYN_OK;
}

/* route the scalar all-reduces and the halo exchanges of this communicator through peer memory (NULL: back to NCCL) */
int pyn_comm_set_peer(pyn_comm* c, pyn_peer* p) {
  PYN_REQUIRE(c, "comm is NULL");
  PYN_REQUIRE(!p || (p->opened && p->rank == c->rank && p->world == c->world), "peer region does not match the communicator");
  c->peer = p;
  return PYN_OK;
}

int pyn_halo_create(const pyn_comm* c, int nnbr, const int32_t* h_nbr_rank, const int64_t* h_send_cnt,
                    const int64_t* const* d_send_idx, const int64_t* h_recv_cnt, const int64_t* const* d_recv_idx,
                    pyn_halo** out) {
  PYN_REQUIRE(c && out && nnbr >= 0 && nnbr <= PYN_HALO_MAX_NBR, "bad arguments");
  pyn_halo* h = new pyn_halo();
  memset(h, 0, sizeof(*h));
  h->comm = c;
  h->nnbr = nnbr;
  for (int k = 0; k < nnbr; ++k) {
    PYN_REQUIRE(h_nbr_rank[k] >= 0 && h_nbr_rank[k] < c->world && h_nbr_rank[k] != c->rank, "bad neighbour rank");
    h->nbr[k] = h_nbr_rank[k];
    h->send_cnt[k] = h_send_cnt[k];
    h->recv_cnt[k] = h_recv_cnt[k];
    h->send_idx[k] = d_send_idx[k];
    h->recv_idx[k] = d_recv_idx[k];
    if (h->send_cnt[k] > 0)
      PYN_CHECK_CUDA(cudaMalloc(&h->send_buf[k], sizeof(double) * h->send_cnt[k] * PYN_HALO_MAX_COMP));
    if (h->recv_cnt[k] > 0)
      PYN_CHECK_CUDA(cudaMalloc(&h->recv_buf[k], sizeof(double) * h->recv_cnt[k] * PYN_HALO_MAX_COMP));
  }
  *out = h;
  return PYN_OK;
}

int pyn_halo_destroy(pyn_halo* h) {
  if (!h) return PYN_OK;
  for (int k = 0; k < h->nnbr; ++k) {
    cudaFree(h->send_buf[k]);
    cudaFree(h->recv_buf[k]);
  }
  delete h;
  return PYN_OK;
}

int pyn_halo_exchange(const pyn_halo* h, double* d_vec, int ncomp, void* stream) {
  PYN_REQUIRE(h && d_vec, "bad arguments");
  return pyn::halo_exchange(h, d_vec, ncomp, pyn::as_stream(stream));
}

int64_t pyn_halo_bytes(const pyn_halo* h, int ncomp) {
  if (!h) return 0;
  int64_t b = 0;
  for (int k = 0; k < h->nnbr; ++k) b += (h->send_cnt[k] + h->recv_cnt[k]) * ncomp * 8;
  return b;
}
}
