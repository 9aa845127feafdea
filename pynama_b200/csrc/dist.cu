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

}  // namespace

namespace pyn {

int comm_allreduce_sum(const pyn_comm* c, double* d_buf, int n, cudaStream_t st) {
  if (!c || c->world == 1) return PYN_OK;
  PYN_CHECK_NCCL(g_nccl.AllReduce(d_buf, d_buf, (size_t)n, ncclFloat64, ncclSum, (ncclComm_t)c->nccl, st));
  return PYN_OK;
}

int halo_exchange(const pyn_halo* h, double* d_vec, int ncomp, cudaStream_t st) {
  if (!h || h->nnbr == 0) return PYN_OK;
  PYN_REQUIRE(ncomp >= 1 && ncomp <= PYN_HALO_MAX_COMP, "halo: ncomp out of range");
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
  c->rank = rank; c->world = world; c->nccl = nullptr;
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
