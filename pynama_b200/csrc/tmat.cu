// Element operator kernels (see tmat.cuh): y = A x with fused epilogues, same modes as pyn_spmv.
#include <stdlib.h>

#include <vector>

#include "tmat.cuh"

namespace {

using pyn::EO_Q;
using pyn::EO_TE;
constexpr int EO_MAXW = 5;   // warps per CTA of phase 1 (<= 160 threads x 3 CTAs per SM caps the registers at 136)

struct ModeEpi {
  double* y;
  int mode;
  double p0, p1;
  const double* aux;
  __device__ __forceinline__ void operator()(int64_t row, double s) const {
    double out;
    switch (mode) {
      case PYN_SPMV_ADD: out = y[row] + s; break;
      case PYN_SPMV_SCALE: out = p0 * s; break;
      case PYN_SPMV_SRT_VTENSV: {
        const int64_t k = row / 3;
        const int c = (int)(row - 3 * k);
        const double vx = aux[2 * k], vy = aux[2 * k + 1];
        const double t = (c == 0) ? vx * vx : ((c == 1) ? vx * vy : vy * vy);
        out = p0 * s + (-p1) * t;
      } break;
      default: out = s;
    }
    y[row] = out;
  }
};

// E (row major [NR][NC]) -> sign * E^T as [NCP][NRP], zero padded
__global__ void eop_transpose_kernel(int NR, int NC, int NRP, int NCP, double sign, const double* __restrict__ E,
                                     double* __restrict__ Et) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < NCP * NRP; i += gridDim.x * blockDim.x) {
    const int k = i / NRP, r = i - k * NRP;
    Et[i] = (r < NR && k < NC) ? sign * E[r * NC + k] : 0.0;
  }
}

// Phase 1.  Persistent CTAs; E^T stays in shared memory for the life of the CTA.  Per tile of 128 elements:
//   gather   threads walk the tile's (element, local node) pairs in memory order: connectivity reads are coalesced,
//            the x loads of neighbouring local nodes fall into the same lines, X lands element-major in shared memory
//   compute  work item = R rows x 4 elements per lane (elements lane, lane+32, ...); per pair of k: 4 LDS.128 of X +
//            R broadcast LDS.128 of E^T feed 8R DFMA.  Items are dealt round-robin to the warps; R and the CTA width
//            are picked per matrix shape so that no warp idles (K at ngl 5: R = 10, 5 warps, 5 items)
//   store    each thread writes its R consecutive rows of 4 elements (runs of 8R bytes of the element-major Ye)
template <int CD, int R>
__global__ void __launch_bounds__(EO_MAXW * 32, 3)
eop_apply_kernel(pyn::TmatView m, const double* __restrict__ x, const int* __restrict__ done) {
  extern __shared__ __align__(16) double sm[];
  if (done && *done) return;
  const int NR = m.NR, NC = m.NC, NRP = m.NRP, NCP = m.NCP, nn = m.nn;
  double* Et = sm;                  // [NCP][NRP]
  double* Xs = sm + NCP * NRP;      // [EO_TE][NCP]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthr = blockDim.x, nwarps = nthr >> 5;
  for (int i = tid; i < NCP * NRP; i += nthr) Et[i] = m.Et[i];
  for (int i = tid; i < EO_TE * (NCP - NC); i += nthr) {   // pad columns of X: zero for the life of the CTA
    const int el = i / (NCP - NC);
    Xs[el * NCP + NC + (i - el * (NCP - NC))] = 0.0;
  }
  const int64_t ntiles = (m.ne + EO_TE - 1) / EO_TE;
  const int nitems = NRP / R;
  constexpr int U = (CD == 3) ? 12 : 20;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t e0 = tile * EO_TE;
    const int64_t rem = m.ne - e0;
    const int npairs = (int)(rem < EO_TE ? rem : EO_TE) * nn;
    const int* en = m.e2n + e0 * nn;
    __syncthreads();   // the previous tile's products are done (and E^T is in place)
    // U independent connectivity loads, then U independent x loads, then the shared-memory stores: two memory round
    // trips per batch instead of two per pair (the gather is latency-bound; the accumulators are not live here)
    for (int base = tid; base < EO_TE * nn; base += U * nthr) {
      int nd[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = base + u * nthr;
        nd[u] = (i < npairs) ? pyn::ld_stream(en + i) : -1;
      }
      double v[U][CD];
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int b = 0; b < CD; ++b) v[u][b] = 0.0;
        if (nd[u] >= 0) {
          if (CD == 2) {
            const double2 t = __ldg(reinterpret_cast<const double2*>(x) + nd[u]);
            v[u][0] = t.x;
            v[u][CD - 1] = t.y;
          } else {
#pragma unroll
            for (int b = 0; b < CD; ++b) v[u][b] = __ldg(x + (int64_t)nd[u] * CD + b);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = base + u * nthr;
        if (i < EO_TE * nn) {
          const int el = (int)__umulhi((unsigned)i, m.nn_magic), j = i - el * nn;
          double* xs = Xs + el * NCP + j * CD;
          if (CD == 2) {
            *reinterpret_cast<double2*>(xs) = make_double2(v[u][0], v[u][CD - 1]);   // NCP and j*2 are even
          } else {
#pragma unroll
            for (int b = 0; b < CD; ++b) xs[b] = v[u][b];
          }
        }
      }
    }
    __syncthreads();
    for (int item = warp; item < nitems; item += nwarps) {
      const int i0 = item * R;
      double acc[R][EO_Q];
#pragma unroll
      for (int r = 0; r < R; ++r)
#pragma unroll
        for (int q = 0; q < EO_Q; ++q) acc[r][q] = 0.0;
      const double* xp = Xs + lane * NCP;
      const double* ep = Et + i0;
      for (int k = 0; k < NCP; k += 2) {
        double2 xv[EO_Q];
#pragma unroll
        for (int q = 0; q < EO_Q; ++q) xv[q] = *reinterpret_cast<const double2*>(xp + q * 32 * NCP + k);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          double ev[R];
#pragma unroll
          for (int r = 0; r < R; r += 2) {
            const double2 t = *reinterpret_cast<const double2*>(ep + (k + kk) * NRP + r);
            ev[r] = t.x;
            ev[r + 1] = t.y;
          }
#pragma unroll
          for (int r = 0; r < R; ++r)
#pragma unroll
            for (int q = 0; q < EO_Q; ++q) acc[r][q] = fma(ev[r], kk ? xv[q].y : xv[q].x, acc[r][q]);
        }
      }
#pragma unroll
      for (int q = 0; q < EO_Q; ++q) {
        const int64_t e = e0 + lane + 32 * q;
        if (e >= m.ne) continue;
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (i0 + r < NR) {
            const int2 rm = __ldg(m.rowmap + i0 + r);
            m.Ye[rm.x + e * rm.y] = acc[r][q];
          }
      }
    }
  }
}

// ---- Phase 1 on the FP64 tensor cores ------------------------------------------------------------------------------
// The SIMT kernel above is co-limited by the FP64 pipe and the shared-memory pipe (every DFMA needs an operand from
// shared memory; ncu: fp64 32 %, LSU 48 %, the rest barrier / gather latency).  This variant computes the same
// Ye = E . X with mma.sync.m8n8k4.f64 (DMMA): A = 8 rows x 4 cols of E (pre-arranged fragment-major in shared memory,
// one conflict-free LDS.64 per fragment), B = 4 cols x 8 elements of X, D = 8 rows x 8 elements -- 256 FMA for two
// operand loads.  Every WARP is autonomous: it owns stages of 8 elements, gathers them asynchronously (cp.async
// straight into its private double buffer, connectivity prefetched one stage ahead into registers) and multiplies
// the previous stage meanwhile, MB row blocks at a time.  No CTA barrier after the prologue: the 24 warps of an SM
// drift apart and keep the DMMA pipe fed while others wait on memory.
constexpr int EM_NBW = 2;     // DMMA n-blocks (8 elements each) per warp stage: every A fragment feeds EM_NBW DMMAs
constexpr int EM_NE = 8 * EM_NBW;   // elements per warp stage
constexpr int EM_PMAX = 16;   // (element, node) pairs gathered per lane and stage, at most (kernel variants: 8, 16)
constexpr int EM_W = 4;       // warps per CTA

__global__ void eop_fragment_kernel(int NR, int NC, int RB, int KB, double sign, const double* __restrict__ E,
                                    double* __restrict__ Af) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < RB * KB * 32; i += gridDim.x * blockDim.x) {
    const int l = i & 31, f = i >> 5, kb = f % KB, rb = f / KB;
    const int r = rb * 8 + (l >> 2), k = kb * 4 + (l & 3);
    Af[i] = (r < NR && k < NC) ? sign * E[r * NC + k] : 0.0;
  }
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int BYTES>
__device__ __forceinline__ void cp_async_zfill(double* smem_dst, const double* gsrc, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  const int n = valid ? BYTES : 0;   // src-size 0: the destination is zero-filled
  asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" ::"r"(d), "l"(gsrc), "n"(BYTES), "r"(n) : "memory");
}

// AG: the A fragments stay in global memory and are read through L1/L2 (element matrices of ngl 8-9: 135-220 kB of
// fragments + the warps' X buffers do not fit in shared memory together; the re-reads cost ne/16 x |A| ~ 1.5 GB of L2
// traffic per product at 16 M dofs, against 30+ GB of the CSR stream these orders used before).
template <int CD, int PMAX, int MB, bool AG = false>
__global__ void __launch_bounds__(EM_W * 32, AG ? 1 : 3)
eop_apply_mma_kernel(pyn::TmatView m, const double* __restrict__ x, const int* __restrict__ done, int RB, int KB) {
  extern __shared__ __align__(16) double sm[];
  if (done && *done) return;
  const int NR = m.NR, NC = m.NC, NCP = m.NCP, nn = m.nn;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double* Af = AG ? m.Et : sm;                              // [RB][KB][32]
  double* Xw = sm + (AG ? 0 : RB * KB * 32) + warp * (2 * EM_NE * NCP);      // this warp's [2][EM_NE][NCP]
  if (!AG)
    for (int i = tid; i < RB * KB * 32; i += EM_W * 32) sm[i] = m.Et[i];
  for (int i = lane; i < 2 * EM_NE * (NCP - NC); i += 32) {        // pad columns of X: zero for the life of the CTA
    const int el = i / (NCP - NC);
    Xw[el * NCP + NC + (i - el * (NCP - NC))] = 0.0;
  }
  __syncthreads();
  const int64_t nstages = (m.ne + EM_NE - 1) / EM_NE;
  const int npt = EM_NE * nn;             // pairs per stage
  int idx[PMAX];
  auto load_idx = [&](int64_t stage) {
    const int64_t e0 = stage * EM_NE;
    const int64_t rem = m.ne - e0;
    const int npairs = rem <= 0 ? 0 : (int)(rem < EM_NE ? rem : EM_NE) * nn;
    const int* en = m.e2n + e0 * nn;
#pragma unroll
    for (int u = 0; u < PMAX; ++u) {
      const int i = lane + u * 32;
      idx[u] = (i < npairs) ? pyn::ld_stream(en + i) : -1;
    }
  };
  auto issue = [&](int buf) {
    double* X = Xw + buf * (EM_NE * NCP);
#pragma unroll
    for (int u = 0; u < PMAX; ++u) {
      const int i = lane + u * 32;
      if (i < npt) {
        const int el = (int)__umulhi((unsigned)i, m.nn_magic), j = i - el * nn;
        double* dst = X + el * NCP + j * CD;
        const bool ok = idx[u] >= 0;
        const double* src = x + (ok ? (int64_t)idx[u] * CD : 0);
        if (CD == 2) {
          cp_async_zfill<16>(dst, src, ok);
        } else {
#pragma unroll
          for (int b = 0; b < CD; ++b) cp_async_zfill<8>(dst + b, src + b, ok);
        }
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int64_t G = (int64_t)gridDim.x * EM_W;
  int64_t stage = (int64_t)blockIdx.x * EM_W + warp;
  load_idx(stage);
  issue(0);
  load_idx(stage + G);
  int buf = 0;
  for (; stage < nstages; stage += G, buf ^= 1) {
    __syncwarp();                         // the products of the previous stage no longer read buf ^ 1
    if (stage + G < nstages) {
      issue(buf ^ 1);                     // stage t+1 streams in while stage t is multiplied
      load_idx(stage + 2 * G);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncwarp();                         // every lane's copies of stage t have landed
    const double* bp = Xw + buf * (EM_NE * NCP) + (lane >> 2) * NCP + (lane & 3);
    const int64_t e0 = stage * EM_NE + (lane & 3) * 2;
    for (int rb0 = 0; rb0 < RB; rb0 += MB) {
      double acc[MB][EM_NBW][2];
#pragma unroll
      for (int i = 0; i < MB; ++i)
#pragma unroll
        for (int nb = 0; nb < EM_NBW; ++nb) acc[i][nb][0] = acc[i][nb][1] = 0.0;
      const double* ap = Af + rb0 * KB * 32 + lane;
#pragma unroll 2
      for (int kb = 0; kb < KB; ++kb) {
        double b[EM_NBW];
#pragma unroll
        for (int nb = 0; nb < EM_NBW; ++nb) b[nb] = bp[nb * 8 * NCP + kb * 4];
        double a[MB];
#pragma unroll
        for (int i = 0; i < MB; ++i) a[i] = (rb0 + i < RB) ? (AG ? __ldg(ap + (i * KB + kb) * 32) : ap[(i * KB + kb) * 32]) : 0.0;
#pragma unroll
        for (int i = 0; i < MB; ++i)
#pragma unroll
          for (int nb = 0; nb < EM_NBW; ++nb) dmma884(acc[i][nb][0], acc[i][nb][1], a[i], b[nb]);
      }
      // row map entries of this pass: MB independent loads issued back to back (one exposed latency per pass; L1 is a
      // few kB next to the shared-memory carve-out and the gathers run through it, so these come from L2).  volatile:
      // re-read per pass instead of living in 2*MB registers across the DMMA loop
      int bx[MB], by[MB];
#pragma unroll
      for (int i = 0; i < MB; ++i) {
        const int row = (rb0 + i) * 8 + (lane >> 2);
        asm volatile("ld.global.nc.v2.s32 {%0,%1}, [%2];" : "=r"(bx[i]), "=r"(by[i]) : "l"(m.rowmap + (row < NR ? row : 0)));
      }
#pragma unroll
      for (int i = 0; i < MB; ++i) {
        const int row = (rb0 + i) * 8 + (lane >> 2);
        if (row < NR) {
          double* yp = m.Ye + bx[i];
#pragma unroll
          for (int nb = 0; nb < EM_NBW; ++nb) {
            const int64_t e = e0 + nb * 8;
            if (e < m.ne) yp[e * by[i]] = acc[i][nb][0];
            if (e + 1 < m.ne) yp[(e + 1) * by[i]] = acc[i][nb][1];
          }
        }
      }
    }
  }
}

// Phase 2: one thread per un-flagged node
template <int RD>
__global__ void __launch_bounds__(256) eop_gather_kernel(pyn::TmatView m, ModeEpi epi) {
  for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < m.nnodes; r += (int64_t)gridDim.x * blockDim.x) {
    if (m.flags[r]) continue;
    double acc[RD];
    pyn::eop_gather_node<RD>(m, r, acc);
#pragma unroll
    for (int a = 0; a < RD; ++a) epi(r * RD + a, acc[a]);
  }
}

// rows of the flagged nodes from the assembled CSR: one warp per row, one row per warp (full-width grid)
__global__ void __launch_bounds__(256)
tmat_boundary_kernel(pyn::TmatView m, int rd, const double* __restrict__ x, ModeEpi epi) {
  const int lane = threadIdx.x & 31;
  const int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  if (i >= m.nlist * rd) return;
  const int64_t row = (int64_t)m.list[i / rd] * rd + (i % rd);
  const double s = pyn::spmv_row<32>(m.ptr, m.ind, m.val, x, row, true, lane);
  if (lane == 0) epi(row, s);
}

// launch shape for an element matrix with NR rows: rows per register block R (8 or 10), warps per CTA W (2..5) and the
// padded sizes; minimises the idle fraction  1 - NR / (R * W * rounds)  (ties: wider CTA, then R = 8)
struct EopShape {
  int R, W, NRP, NCP;
  size_t smem;
};
EopShape eop_shape(int NR, int NC) {
  EopShape best{8, 4, 0, 0, 0};
  double best_eff = -1.0;
  for (int R : {8, 10}) {
    const int items = (NR + R - 1) / R;
    for (int W = EO_MAXW; W >= 2; --W) {
      const int rounds = (items + W - 1) / W;
      const double eff = (double)NR / ((double)R * W * rounds);
      if (eff > best_eff + 1e-9) { best_eff = eff; best.R = R; best.W = W; }
    }
  }
  best.NRP = (NR + best.R - 1) / best.R * best.R;
  int ncp = (NC + 1) / 2 * 2;          // even, and NCP/2 odd: the LDS.128 reads of X are then bank-conflict free
  if ((ncp / 2) % 2 == 0) ncp += 2;
  best.NCP = ncp;
  best.smem = sizeof(double) * ((size_t)best.NCP * best.NRP + (size_t)EO_TE * best.NCP);
  return best;
}
constexpr size_t EO_SMEM_MAX = 227 * 1024;

// DMMA variant: row blocks, k blocks, padded columns (NCP = 4 mod 8: conflict-free B fragments), row blocks per pass
// MB (3..7: least padding, then the larger), shared memory
struct MmaShape {
  int RB, KB, NCP, MB, P;
  size_t smem;
  bool ok;
  bool ag;     // A fragments from L1/L2 (they do not fit in shared memory next to the X buffers)
};
constexpr int EM_PMAX_AG = 48;
MmaShape mma_shape(int nn, int NR, int NC) {
  MmaShape s;
  s.RB = (NR + 7) / 8;
  int ncp = (NC + 3) / 4 * 4;
  if (ncp % 8 == 0) ncp += 4;
  s.NCP = ncp; s.KB = ncp / 4;
  s.MB = 3;
  int best = 1 << 30;
  for (int MB = 7; MB >= 3; --MB) {
    const int waste = (s.RB + MB - 1) / MB * MB - s.RB;
    if (waste < best) { best = waste; s.MB = MB; }
  }
  s.P = (EM_NE * nn + 31) / 32;
  s.smem = sizeof(double) * ((size_t)s.RB * s.KB * 32 + (size_t)EM_W * 2 * EM_NE * s.NCP);
  s.ok = s.smem <= EO_SMEM_MAX && s.P <= EM_PMAX;
  s.ag = false;
  if (!s.ok) {   // large element matrices: X buffers only in shared memory, MB in {4, 7}
    const size_t xs = sizeof(double) * (size_t)EM_W * 2 * EM_NE * s.NCP;
    if (xs <= EO_SMEM_MAX && s.P <= EM_PMAX_AG) {
      s.ok = s.ag = true;
      s.smem = xs;
      const int w4 = (s.RB + 3) / 4 * 4 - s.RB, w7 = (s.RB + 6) / 7 * 7 - s.RB;
      s.MB = w7 <= w4 ? 7 : 4;
    }
  }
  return s;
}
template <int CD, int PMAX, int MB, bool AG = false>
int mma_launch(const pyn_tmat* m, const double* d_x, const int* d_done, cudaStream_t st) {
  eop_apply_mma_kernel<CD, PMAX, MB, AG><<<m->grid_apply, 32 * EM_W, m->smem, st>>>(pyn::view_of(m), d_x, d_done, m->RB,
                                                                                     m->KB);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}
template <int CD, int PMAX, int MB, bool AG = false>
int mma_prepare(pyn_tmat* m) {
  int occ = 0;
  PYN_CHECK_CUDA(cudaFuncSetAttribute(eop_apply_mma_kernel<CD, PMAX, MB, AG>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)EO_SMEM_MAX));
  if (!AG)   // the AG variant keeps what shared memory it does not need as L1 for the A fragments
    PYN_CHECK_CUDA(cudaFuncSetAttribute(eop_apply_mma_kernel<CD, PMAX, MB, AG>,
                                        cudaFuncAttributePreferredSharedMemoryCarveout,
                                        (int)cudaSharedmemCarveoutMaxShared));
  PYN_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, eop_apply_mma_kernel<CD, PMAX, MB, AG>, 32 * EM_W,
                                                               m->smem));
  const int64_t nstages = (m->topo->ne + EM_NE - 1) / EM_NE;
  const int64_t need = (nstages + EM_W - 1) / EM_W;
  const int64_t cap = (int64_t)pyn::sm_count() * (occ > 0 ? occ : 1);
  m->grid_apply = (int)(need < cap ? need : cap);
  return PYN_OK;
}

template <int CD, int R>
int eop_launch(const pyn_tmat* m, const double* d_x, const int* d_done, cudaStream_t st) {
  eop_apply_kernel<CD, R><<<m->grid_apply, 32 * m->W, m->smem, st>>>(pyn::view_of(m), d_x, d_done);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}
template <int CD, int R>
int eop_prepare(pyn_tmat* m) {
  int occ = 0;
  PYN_CHECK_CUDA(cudaFuncSetAttribute(eop_apply_kernel<CD, R>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)EO_SMEM_MAX));
  if (!getenv("PYN_EO_NO_CARVEOUT"))
    PYN_CHECK_CUDA(cudaFuncSetAttribute(eop_apply_kernel<CD, R>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                        (int)cudaSharedmemCarveoutMaxShared));
  PYN_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, eop_apply_kernel<CD, R>, 32 * m->W, m->smem));
  const int64_t ntiles = (m->topo->ne + EO_TE - 1) / EO_TE;
  const int64_t cap = (int64_t)pyn::sm_count() * (occ > 0 ? occ : 1);
  m->grid_apply = (int)(ntiles < cap ? ntiles : cap);
  return PYN_OK;
}
#define PYN_EO_DISPATCH(FN, ...)                                                     \
  do {                                                                               \
    if (m->R == 8) {                                                                 \
      if (m->cdof == 1) return FN<1, 8>(__VA_ARGS__);                                \
      if (m->cdof == 2) return FN<2, 8>(__VA_ARGS__);                                \
      return FN<3, 8>(__VA_ARGS__);                                                  \
    }                                                                                \
    if (m->cdof == 1) return FN<1, 10>(__VA_ARGS__);                                 \
    if (m->cdof == 2) return FN<2, 10>(__VA_ARGS__);                                 \
    return FN<3, 10>(__VA_ARGS__);                                                   \
  } while (0)
int eop_prepare_any(pyn_tmat* m) { PYN_EO_DISPATCH(eop_prepare, m); }
// DMMA variant: PMAX = 8 when a lane gathers at most 8 pairs per stage (fewer registers), else 16; MB = 3..7
#define PYN_EM_CASE(FN, CDV, PV, ...)                                                \
  switch (m->MB) {                                                                   \
    case 3: return FN<CDV, PV, 3>(__VA_ARGS__);                                      \
    case 4: return FN<CDV, PV, 4>(__VA_ARGS__);                                      \
    case 5: return FN<CDV, PV, 5>(__VA_ARGS__);                                      \
    case 6: return FN<CDV, PV, 6>(__VA_ARGS__);                                      \
    default: return FN<CDV, PV, 7>(__VA_ARGS__);                                     \
  }
#define PYN_EM_AG_CASE(FN, CDV, ...)                                                 \
  if (m->P <= 32) {                                                                  \
    if (m->MB == 4) return FN<CDV, 32, 4, true>(__VA_ARGS__);                        \
    return FN<CDV, 32, 7, true>(__VA_ARGS__);                                        \
  } else {                                                                           \
    if (m->MB == 4) return FN<CDV, 48, 4, true>(__VA_ARGS__);                        \
    return FN<CDV, 48, 7, true>(__VA_ARGS__);                                        \
  }
#define PYN_EM_DISPATCH(FN, ...)                                                     \
  do {                                                                               \
    if (m->ag) {                                                                     \
      if (m->cdof == 1) { PYN_EM_AG_CASE(FN, 1, __VA_ARGS__) }                       \
      if (m->cdof == 2) { PYN_EM_AG_CASE(FN, 2, __VA_ARGS__) }                       \
      PYN_EM_AG_CASE(FN, 3, __VA_ARGS__)                                             \
    }                                                                                \
    if (m->P <= 8) {                                                                 \
      if (m->cdof == 1) { PYN_EM_CASE(FN, 1, 8, __VA_ARGS__) }                       \
      if (m->cdof == 2) { PYN_EM_CASE(FN, 2, 8, __VA_ARGS__) }                       \
      PYN_EM_CASE(FN, 3, 8, __VA_ARGS__)                                             \
    }                                                                                \
    if (m->cdof == 1) { PYN_EM_CASE(FN, 1, 16, __VA_ARGS__) }                        \
    if (m->cdof == 2) { PYN_EM_CASE(FN, 2, 16, __VA_ARGS__) }                        \
    PYN_EM_CASE(FN, 3, 16, __VA_ARGS__)                                              \
  } while (0)
int mma_prepare_any(pyn_tmat* m) { PYN_EM_DISPATCH(mma_prepare, m); }

}  // namespace

namespace pyn {

int tmat_apply(const pyn_tmat* m, const double* d_x, const int* d_done, cudaStream_t st) {
  if (m->mma) PYN_EM_DISPATCH(mma_launch, m, d_x, d_done, st);
  PYN_EO_DISPATCH(eop_launch, m, d_x, d_done, st);
}

// called by pyn_tmat_create (assemble.cu, which owns the descriptor logic): element matrix, scratch, launch shapes
int tmat_init(pyn_tmat* m, const double* d_E, double sign, cudaStream_t st) {
  const pyn_topo* t = m->topo;
  m->NR = m->rdof * t->nn; m->NC = m->cdof * t->nn;
  const MmaShape ms = mma_shape(t->nn, m->NR, m->NC);
  // the L2-fed DMMA variant only where the DFMA kernel does not fit either (ngl 8-9); PYN_EO_SIMT=1: A/B measurements
  const bool simt_fits = eop_shape(m->NR, m->NC).smem <= EO_SMEM_MAX;
  m->mma = ms.ok && (!ms.ag || !simt_fits) && !(getenv("PYN_EO_SIMT") && simt_fits);
  m->ag = m->mma && ms.ag;
  {
    // Row map of Ye: one dense array per node kind for square 2-D elements in the reference's local order (4 vertices,
    // 4 x (n-2) edge nodes, (n-2)^2 interior nodes); anything else, or PYN_EO_AOS=1, keeps element-major records.
    const int nn = t->nn, rd = m->rdof;
    int n = 1;
    while (n * n < nn) ++n;
    std::vector<int> first;   // first local node of every kind, then nn
    if (n * n == nn && n >= 3 && !getenv("PYN_EO_AOS") && (int64_t)t->ne * m->NR + 64 * 8 < 0x7fffffff) {
      first.push_back(0);
      for (int k = 0; k <= 4; ++k) first.push_back(4 + k * (n - 2));
    } else {
      first.push_back(0);
    }
    first.push_back(nn);
    std::vector<int2> map((size_t)m->NR);
    int64_t base = 0;
    for (size_t g = 0; g + 1 < first.size(); ++g) {
      const int stride = (first[g + 1] - first[g]) * rd;
      for (int l = first[g]; l < first[g + 1]; ++l)
        for (int a = 0; a < rd; ++a) map[(size_t)l * rd + a] = make_int2((int)base + (l - first[g]) * rd + a, stride);
      base += (int64_t)t->ne * stride;
      base = (base + 15) & ~(int64_t)15;     // every array starts on a 128-byte line
    }
    PYN_REQUIRE(first.size() == 2 || base < 0x7fffffff, "element operator: row map overflow");
    PYN_CHECK_CUDA(cudaMalloc(&m->Ye, sizeof(double) * (size_t)base));
    PYN_CHECK_CUDA(cudaMalloc(&m->rowmap, sizeof(int2) * map.size()));
    PYN_CHECK_CUDA(cudaMemcpyAsync(m->rowmap, map.data(), sizeof(int2) * map.size(), cudaMemcpyHostToDevice, st));
    PYN_CHECK_CUDA(cudaStreamSynchronize(st));   // `map` is pageable host memory going out of scope
  }
  int rc;
  if (m->mma) {
    m->RB = ms.RB; m->KB = ms.KB; m->NCP = ms.NCP; m->W = EM_W; m->MB = ms.MB; m->P = ms.P; m->smem = ms.smem;
    m->NRP = ms.RB * 8;
    const int len = m->RB * m->KB * 32;
    PYN_CHECK_CUDA(cudaMalloc(&m->Et, sizeof(double) * (size_t)len));
    eop_fragment_kernel<<<(len + 255) / 256, 256, 0, st>>>(m->NR, m->NC, m->RB, m->KB, sign, d_E, m->Et);
    PYN_CHECK_LAUNCH();
    rc = mma_prepare_any(m);
  } else {
    const EopShape sh = eop_shape(m->NR, m->NC);
    m->R = sh.R; m->W = sh.W; m->NRP = sh.NRP; m->NCP = sh.NCP; m->smem = sh.smem;
    PYN_REQUIRE(m->smem <= EO_SMEM_MAX, "element operator: the element matrix does not fit in shared memory");
    PYN_CHECK_CUDA(cudaMalloc(&m->Et, sizeof(double) * (size_t)m->NCP * m->NRP));
    eop_transpose_kernel<<<(m->NCP * m->NRP + 255) / 256, 256, 0, st>>>(m->NR, m->NC, m->NRP, m->NCP, sign, d_E, m->Et);
    PYN_CHECK_LAUNCH();
    rc = eop_prepare_any(m);
  }
  if (rc) return rc;
  m->grid_int = grid_for(t->nnodes, 256, 8);
  const int64_t need_b = (m->nlist * m->rdof * 32 + 255) / 256;
  PYN_REQUIRE(need_b < 0x7fffffff, "element operator: boundary list too long");
  m->grid_bnd = (int)(need_b > 0 ? need_b : 1);
  PYN_CHECK_CUDA(cudaStreamSynchronize(st));
  return PYN_OK;
}

}  // namespace pyn

extern "C" {

int pyn_tmat_supported(int nn, int rdof, int cdof) {
  if (nn < 1 || rdof < 1 || rdof > 3 || cdof < 1 || cdof > 3) return 0;
  return (mma_shape(nn, rdof * nn, cdof * nn).ok || eop_shape(rdof * nn, cdof * nn).smem <= EO_SMEM_MAX) ? 1 : 0;
}

int pyn_tmat_destroy(pyn_tmat* m) {
  if (!m) return PYN_OK;
  cudaFree(m->Et);
  cudaFree(m->Ye);
  cudaFree(m->rowmap);
  delete m;
  return PYN_OK;
}

int pyn_tmat_mult(const pyn_tmat* m, const double* d_x, double* d_y, int mode, double p0, double p1, const double* d_aux,
                  void* stream) {
  PYN_REQUIRE(m && d_x && d_y, "bad arguments");
  PYN_REQUIRE(mode >= 0 && mode <= 3, "bad mode");
  PYN_REQUIRE(mode != PYN_SPMV_SRT_VTENSV || (d_aux != nullptr && m->rdof == 3), "SRT_VTENSV needs aux and 3 rows per node");
  cudaStream_t st = pyn::as_stream(stream);
  const pyn::TmatView v = pyn::view_of(m);
  ModeEpi epi{d_y, mode, p0, p1, d_aux};
  int rc = pyn::tmat_apply(m, d_x, nullptr, st);
  if (rc) return rc;
  if (m->nlist > 0) {
    tmat_boundary_kernel<<<m->grid_bnd, 256, 0, st>>>(v, m->rdof, d_x, epi);
    PYN_CHECK_LAUNCH();
  }
  if (m->rdof == 1) eop_gather_kernel<1><<<m->grid_int, 256, 0, st>>>(v, epi);
  else if (m->rdof == 2) eop_gather_kernel<2><<<m->grid_int, 256, 0, st>>>(v, epi);
  else eop_gather_kernel<3><<<m->grid_int, 256, 0, st>>>(v, epi);
  PYN_CHECK_LAUNCH();
  return PYN_OK;
}

/* algorithmic HBM bytes of one product (tmat.cuh): x once, connectivity, Ye written and read once, the incidence
 * lists, y once; plus the CSR rows of the flagged nodes at 12 B per entry (not counted: O(boundary)) */
int64_t pyn_tmat_bytes(const pyn_tmat* m) {
  if (!m) return 0;
  const pyn_topo* t = m->topo;
  const int64_t inc = t->ne * (int64_t)t->nn;
  return 8 * t->nnodes * m->cdof + 4 * inc + 16 * t->ne * (int64_t)m->NR + 8 * (t->nnodes + 1) + 6 * inc +
         8 * t->nnodes * m->rdof;
}
}
