// Communicator + halo-exchange plan shared by dist.cu and ksp.cu.
#pragma once
#include "common.cuh"

#define PYN_HALO_MAX_NBR 4
#define PYN_HALO_MAX_COMP 3

struct pyn_comm {
  int rank, world;
  void* nccl;  // ncclComm_t (NULL when world == 1)
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
int halo_exchange(const pyn_halo* h, double* d_vec, int ncomp, cudaStream_t st);
}  // namespace pyn
