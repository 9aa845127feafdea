// Communicator + halo-exchange plan shared by dist.cu and ksp.cu.
#pragma once
#include "common.cuh"

#define PYN_HALO_MAX_NBR 4
#define PYN_HALO_MAX_COMP 3

#define PYN_PEER_MAX 8
#define PYN_PEER_SLOT 8   // doubles per all-reduce slot: up to 7 values + the epoch flag

// Peer memory (NVLink P2P through CUDA IPC): every rank owns one region that all the other ranks of the box map.
//   [all-reduce slots]  2 parities x world ranks x PYN_PEER_SLOT doubles: rank r stores its partial sums into slot r of
//                       EVERY rank's region, then the epoch; each rank sums the slots in rank order (deterministic, the
//                       same bits everywhere) -- replaces ncclAllReduce of 1-3 scalars, whose cost is pure latency
//   [halo mailboxes]    2 sides (from the rank below / above) x 2 parities x (8 + cap) doubles: the neighbour's pack
//                       kernel writes the interface values straight into the mailbox and releases an epoch flag
// Parity double-buffering is enough: a rank can only be one exchange ahead of its neighbours (dist.cu).
struct pyn_peer {
  int rank, world;
  double* local;                    // cudaMalloc'd, exported with cudaIpcGetMemHandle
  double* peer[PYN_PEER_MAX];       // every rank's region as seen from here (peer[rank] == local)
  int64_t cap;                      // payload doubles per mailbox
  size_t bytes;
  unsigned long long* ctr;          // device counters: [0] all-reduce epoch, [1] halo epoch (put side), [2] (wait side)
  int opened;
};

struct pyn_comm {
  int rank, world;
  void* nccl;  // ncclComm_t (NULL when world == 1)
  pyn_peer* peer;   // borrowed; NULL = NCCL for everything
};

struct pyn_halo {
  const pyn_comm* comm;
  int nnbr;
  int nbr[PYN_HALO_MAX_NBR];
  int64_t send_cnt[PYN_HALO_MAX_NBR], recv_cnt[PYN_HALO_MAX_NBR];   // in nodes
  const int64_t* send_idx[PYN_HALO_MAX_NBR];                        // borrowed device lists of local node ids
  const int64_t* recv_idx[PYN_HALO_MAX_NBR];
  double* send_buf[PYN_HALO_MAX_NBR];
  double* recv_buf[PYN_HALO_MAX_NBR];
};

namespace pyn {
int comm_allreduce_sum(const pyn_comm* c, double* d_buf, int n, cudaStream_t st);
// out[0] = sum over ranks of sum(pa[0..na)), out[1] = the same for pb (nullable): the per-CTA partials are reduced and
// exchanged by ONE one-warp kernel on the peer path (reduce kernel + ncclAllReduce otherwise)
int comm_reduce_allreduce(const pyn_comm* c, int na, const double* pa, int nb, const double* pb, double* d_out,
                          cudaStream_t st);
int halo_exchange(const pyn_halo* h, double* d_vec, int ncomp, cudaStream_t st);
}  // namespace pyn
