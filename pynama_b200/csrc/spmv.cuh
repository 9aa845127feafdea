// CSR SpMV building block shared by spmv.cu and ksp.cu.
//
// "CSR-vector": a sub-warp of LPR lanes owns one row; lanes stride through the row so that the value / index
// streams are read with coalesced, L1-bypassing loads, the gathered vector goes through the normal cached path,
// and the row sum is finished with a fixed shuffle tree (deterministic).  Rows are distributed grid-stride so the
// number of CTAs (and of per-CTA dot-product partials) is bounded by the SM count.
#pragma once
#include "common.cuh"

namespace pyn {

template <int LPR>
__device__ __forceinline__ double spmv_row(const int64_t* __restrict__ ptr, const int* __restrict__ ind,
                                           const double* __restrict__ val, const double* __restrict__ x, int64_t row,
                                           bool active, int sub) {
  double s0 = 0.0, s1 = 0.0;
  if (active) {
    const int64_t b = ptr[row], e = ptr[row + 1];
    int64_t i = b + sub;
    for (; i + LPR < e; i += 2 * LPR) {  // two independent loads in flight per lane
      const double v0 = ld_stream(val + i), v1 = ld_stream(val + i + LPR);
      const int c0 = ld_stream(ind + i), c1 = ld_stream(ind + i + LPR);
      s0 += v0 * x[c0];
      s1 += v1 * x[c1];
    }
    if (i < e) s0 += ld_stream(val + i) * x[ld_stream(ind + i)];
  }
  return subwarp_sum<LPR>(s0 + s1);
}

static inline int pick_lpr(int64_t nrows, int64_t nnz) {
  const double avg = nrows > 0 ? (double)nnz / (double)nrows : 0.0;
  if (avg <= 6.0) return 4;
  if (avg <= 12.0) return 8;
  if (avg <= 40.0) return 16;
  return 32;
}

}  // namespace pyn

// ---------------------------------------------------------------------------------------------------------------
// "CSR-stream" SpMV: the matrix is cut into chunks of ~SP_C consecutive non-zeros snapped to row boundaries (plan
// built once per matrix).  A CTA streams one chunk's values and column indices with fully coalesced, independent
// loads (8 in flight per thread), multiplies by the gathered x and parks the products in shared memory; the row
// sums are then taken from shared memory by sub-warps whose width adapts to the number of rows in the chunk.  HBM
// efficiency therefore does not depend on the row lengths (1 ... several hundred in these matrices), and the
// summation order is a pure function of the sparsity pattern (bitwise reproducible).
// ---------------------------------------------------------------------------------------------------------------
namespace pyn {

constexpr int SP_BLK = 256;      // threads per CTA
constexpr int SP_U = 8;          // independent (value, index) loads in flight per thread
constexpr int SP_CAP = 3072;     // products held in shared memory per chunk
constexpr int SP_RCAP = 512;     // row offsets staged in shared memory per chunk

struct SpmvPlanView {
  int64_t nrows, nnz;
  const int64_t* ptr;
  const int* blk_row;  // [nblk+1] first row of each chunk
  int nblk;
  // bulk-copy kernels: their own, finer chunk table (smaller stages -> more resident CTAs)
  const int* brow;           // [nbblk+1]
  const long long* bp0;      // [nbblk+1] ptr[brow[b]]: first non-zero of each chunk
  int nbblk;
};

struct SpmvSmem {
  double prod[SP_CAP];
  int off[SP_RCAP + 1];
};

// Processes chunk `b`; calls epi(row, sum) once per row from exactly one thread.
template <class Epi>
__device__ __forceinline__ void spmv_stream_rows(const SpmvPlanView& P, const int* __restrict__ ind,
                                                 const double* __restrict__ val, const double* __restrict__ x, int r0,
                                                 int r1, SpmvSmem& sm, Epi&& epi);
template <class Epi>
__device__ __forceinline__ void spmv_stream_chunk(const SpmvPlanView& P, const int* __restrict__ ind,
                                                  const double* __restrict__ val, const double* __restrict__ x, int b,
                                                  SpmvSmem& sm, Epi&& epi) {
  spmv_stream_rows(P, ind, val, x, P.blk_row[b], P.blk_row[b + 1], sm, epi);
}
// rows [r0, r1) as one chunk
template <class Epi>
__device__ __forceinline__ void spmv_stream_rows(const SpmvPlanView& P, const int* __restrict__ ind,
                                                 const double* __restrict__ val, const double* __restrict__ x, int r0,
                                                 int r1, SpmvSmem& sm, Epi&& epi) {
  const int tid = threadIdx.x;
  const int nr = r1 - r0;
  if (nr <= 0) return;
  const int64_t p0 = P.ptr[r0];
  const int cnt = (int)min(P.ptr[r1] - p0, (int64_t)0x7fffffff);
  if (cnt <= SP_CAP) {
    const bool staged = nr <= SP_RCAP;
    if (staged)
      for (int i = tid; i <= nr; i += SP_BLK) sm.off[i] = (int)(P.ptr[r0 + i] - p0);
    const double* v0 = val + p0;
    const int* c0 = ind + p0;
    for (int base = 0; base < cnt; base += SP_U * SP_BLK) {
      double vv[SP_U];
      int cc[SP_U];
#pragma unroll
      for (int u = 0; u < SP_U; ++u) {
        const int i = base + u * SP_BLK + tid;
        const bool ok = i < cnt;
        vv[u] = ok ? ld_stream(v0 + i) : 0.0;
        cc[u] = ok ? ld_stream(c0 + i) : -1;
      }
#pragma unroll
      for (int u = 0; u < SP_U; ++u) {
        const int i = base + u * SP_BLK + tid;
        if (cc[u] >= 0) sm.prod[i] = vv[u] * x[cc[u]];
      }
    }
    __syncthreads();
    int lpr = 32;
    while (lpr > 1 && lpr * nr > SP_BLK) lpr >>= 1;
    const int rpp = SP_BLK / lpr, sub = tid & (lpr - 1), slot = tid / lpr;
    for (int rb = 0; rb < nr; rb += rpp) {
      const int i = rb + slot;
      const bool active = i < nr;
      double s = 0.0;
      if (active) {
        const int lo = staged ? sm.off[i] : (int)(P.ptr[r0 + i] - p0);
        const int hi = staged ? sm.off[i + 1] : (int)(P.ptr[r0 + i + 1] - p0);
        for (int j = lo + sub; j < hi; j += lpr) s += sm.prod[j];
      }
      for (int o = lpr >> 1; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o, lpr);
      if (active && sub == 0) epi((int64_t)r0 + i, s);
    }
    __syncthreads();
  } else {
    // a chunk holding a row longer than the shared-memory budget: one warp per row straight from global memory
    const int lane = tid & 31, w = tid >> 5;
    for (int i = w; i < nr; i += SP_BLK / 32) {
      const double s = spmv_row<32>(P.ptr, ind, val, x, (int64_t)r0 + i, true, lane);
      if (lane == 0) epi((int64_t)r0 + i, s);
    }
  }
}

}  // namespace pyn

// ---------------------------------------------------------------------------------------------------------------
// "CSR-stream, bulk copy": the same products and row sums as spmv_stream_chunk on a finer chunk table, but the value /
// index / row-pointer slices of a chunk are brought into shared memory by the TMA unit --
// 1-D `cp.async.bulk.shared::cluster.global` completing on an mbarrier (UBLKCP in SASS) -- issued by one thread
// SB_STAGES-1 chunks ahead of the consumers.  The bytes in flight per SM no longer depend on registers or on resident
// warps (two CTAs keep ~100 kB on the wire), which is what bounded the register-staged kernel at 0.71 of the HBM peak
// (DESIGN.md section 3).  Products overwrite the staged values in place; the stage is handed back to the producer
// by the CTA barrier that already separates the row sums of consecutive chunks.
// ---------------------------------------------------------------------------------------------------------------
namespace pyn {

constexpr int SB_BLK = 256;
constexpr int SB_STAGES = 4;
constexpr int SB_CHUNK = 1024;    // non-zeros per bulk chunk (rows snapped: a chunk holds <= SB_CHUNK when rows are shorter)
constexpr int SB_CAP = SB_CHUNK + 16;   // per stage
constexpr int SB_RCAP = 124;      // row pointers staged per chunk (more rows in a chunk: read from global)
constexpr int SB_U = 4;           // x gathers in flight per lane

struct SbStage {
  double val[SB_CAP + 2];               // 16-byte aligned superset of the chunk's values; products in place
  int ind[SB_CAP + 4];
  long long ptr[SB_RCAP + 4];           // row pointers r0 .. r0+nr (aligned superset)
};
struct SbSmem {
  SbStage st[SB_STAGES];
  unsigned long long full[SB_STAGES];
  unsigned long long empty[SB_STAGES];   // mbarriers: bytes landed / all consumer warps done
  int r0[SB_STAGES], nr[SB_STAGES], cnt[SB_STAGES], vo[SB_STAGES], io[SB_STAGES], po[SB_STAGES];
  long long p0[SB_STAGES];
};

__device__ __forceinline__ unsigned smem_addr(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  do {
    asm volatile(
        "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
        : "=r"(ok)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// producer side (one thread): describe chunk b in stage s and start its three bulk copies
__device__ __forceinline__ void sb_issue(const SpmvPlanView& P, const int* __restrict__ ind, const double* __restrict__ val,
                                         int r0, int r1, long long p0, long long p1, SbSmem& sm, int s) {
  const int nr = r1 - r0, cnt = (int)(p1 - p0);
  sm.r0[s] = r0; sm.nr[s] = nr; sm.cnt[s] = cnt; sm.p0[s] = p0;
  unsigned bytes = 0;
  if (nr > 0 && cnt > 0) {
    // aligned supersets: 16-byte boundaries of the global arrays (all three are 16-byte aligned allocations); the
    // over-read at the upper end stays inside the arrays except in the last chunk, which the caller keeps off this path
    const long long va = p0 & ~1ll, ve = (p1 + 1) & ~1ll;
    const long long ia = p0 & ~3ll, ie = (p1 + 3) & ~3ll;
    sm.vo[s] = (int)(p0 - va); sm.io[s] = (int)(p0 - ia);
    const unsigned vb = (unsigned)(ve - va) * 8u, ib = (unsigned)(ie - ia) * 4u;
    unsigned pb = 0;
    if (nr <= SB_RCAP) {
      const long long ra = r0 & ~1ll, re = ((long long)r1 + 2) & ~1ll;   // rows r0 .. r1 inclusive
      sm.po[s] = (int)(r0 - ra);
      pb = (unsigned)(re - ra) * 8u;
    }
    bytes = vb + ib + pb;
    mbar_expect_tx(&sm.full[s], bytes);
    bulk_g2s(sm.st[s].val, val + va, vb, &sm.full[s]);
    bulk_g2s(sm.st[s].ind, ind + ia, ib, &sm.full[s]);
    if (pb) bulk_g2s(sm.st[s].ptr, P.ptr + (r0 & ~1ll), pb, &sm.full[s]);
  } else {
    mbar_expect_tx(&sm.full[s], 0);
  }
}

// All chunks of this CTA (b = first, first + stride, ...).  epi(row, sum) once per row from exactly one thread.
// Requires: every chunk holds <= SB_CAP non-zeros (plan->bulk_ok) and the last chunk of the matrix is excluded
// (nblk_bulk = nblk - 1: its aligned over-read could leave the arrays).
//
// SB_BLK consumer threads + one producer warp (launch with SB_THREADS).  The pipeline is warp-asynchronous: a consumer
// warp waits for a stage to be full, takes an eighth of the chunk's rows -- a contiguous range of entries --, forms the
// products of that range in place (every x gather independent, SB_U per lane in flight), sums its rows by sub-warps
// (fixed shuffle tree: deterministic; the order differs from spmv_stream_chunk's because the chunks do), arrives
// on the stage's `empty` barrier and moves on; the producer refills a stage as soon as all eight warps have left it.
// No CTA-wide barrier inside the loop.
constexpr int SB_THREADS = SB_BLK + 32;

template <class Epi>
__device__ __forceinline__ void spmv_bulk_loop(const SpmvPlanView& P, const int* __restrict__ ind,
                                               const double* __restrict__ val, const double* __restrict__ x, int first,
                                               int stride, int nblk_bulk, SbSmem& sm, Epi&& epi) {
  const int tid = threadIdx.x;
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < SB_STAGES; ++s) {
      mbar_init(&sm.full[s], 1);
      mbar_init(&sm.empty[s], SB_BLK / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nmine = first < nblk_bulk ? (nblk_bulk - first + stride - 1) / stride : 0;
  if (tid >= SB_BLK) {   // ---- producer warp (one lane issues) ----
    if (tid == SB_BLK && nmine > 0) {
      // the chunk descriptors are fetched one chunk ahead, so that their latency is not added to every issue
      int r0 = P.brow[first], r1 = P.brow[first + 1];
      long long p0 = P.bp0[first], p1 = P.bp0[first + 1];
      for (int i = 0; i < nmine; ++i) {
        const int s = i % SB_STAGES;
        const int bn = first + (i + 1 < nmine ? i + 1 : i) * stride;
        const int nr0 = P.brow[bn], nr1 = P.brow[bn + 1];
        const long long np0 = P.bp0[bn], np1 = P.bp0[bn + 1];
        if (i >= SB_STAGES) mbar_wait(&sm.empty[s], (unsigned)(((i / SB_STAGES) - 1) & 1));
        sb_issue(P, ind, val, r0, r1, p0, p1, sm, s);
        r0 = nr0; r1 = nr1; p0 = np0; p1 = np1;
      }
    }
    return;
  }
  const int lane = tid & 31, w = tid >> 5;
  constexpr int NW = SB_BLK / 32;
  for (int i = 0; i < nmine; ++i) {
    const int s = i % SB_STAGES;
    mbar_wait(&sm.full[s], (unsigned)((i / SB_STAGES) & 1));
    const int nr = sm.nr[s], cnt = sm.cnt[s], r0 = sm.r0[s];
    if (nr > 0 && cnt > 0) {
      double* pv = sm.st[s].val + sm.vo[s];
      const int* pi = sm.st[s].ind + sm.io[s];
      const bool staged = nr <= SB_RCAP;
      const long long* pp = sm.st[s].ptr + sm.po[s];
      const long long p0 = sm.p0[s];
      // this warp's rows [ra, rb) and their (contiguous) entries [ea, eb)
      const int ra = (w * nr) / NW, rb = ((w + 1) * nr) / NW;
      if (rb > ra) {
        const int ea = (int)((staged ? pp[ra] : P.ptr[r0 + ra]) - p0), eb = (int)((staged ? pp[rb] : P.ptr[r0 + rb]) - p0);
        // products in place: every gather of the warp's range is independent, SP_U per lane in flight
        for (int base = ea; base < eb; base += SB_U * 32) {
          double vv[SB_U];
          int cc[SB_U];
#pragma unroll
          for (int u = 0; u < SB_U; ++u) {
            const int k = base + u * 32 + lane;
            const bool ok = k < eb;
            cc[u] = ok ? pi[k] : -1;
            vv[u] = ok ? pv[k] : 0.0;
          }
#pragma unroll
          for (int u = 0; u < SB_U; ++u)
            if (cc[u] >= 0) vv[u] = __dmul_rn(vv[u], x[cc[u]]);
#pragma unroll
          for (int u = 0; u < SB_U; ++u)
            if (cc[u] >= 0) pv[base + u * 32 + lane] = vv[u];
        }
        __syncwarp();
        // row sums by sub-warps whose width adapts to the rows in the chunk (a pure function of the pattern)
        int lpr = 32;
        while (lpr > 1 && lpr * nr > SB_BLK) lpr >>= 1;
        const int rpp = 32 / lpr, sub = lane & (lpr - 1), slot = lane / lpr;
        for (int rr = ra; rr < rb; rr += rpp) {
          const int r = rr + slot;
          const bool active = r < rb;
          double acc = 0.0;
          if (active) {
            const int lo = (int)((staged ? pp[r] : P.ptr[r0 + r]) - p0);
            const int hi = (int)((staged ? pp[r + 1] : P.ptr[r0 + r + 1]) - p0);
            for (int j = lo + sub; j < hi; j += lpr) acc += pv[j];
          }
          for (int o = lpr >> 1; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o, lpr);
          if (active && sub == 0) epi((int64_t)r0 + r, acc);
        }
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&sm.empty[s]);   // this warp is done with stage s
  }
}

}  // namespace pyn

struct pyn_spmv_plan {
  int64_t nrows, nnz;
  const int64_t* ptr;  // borrowed
  int* blk_row;
  int* brow;
  long long* bp0;
  int nblk, chunk, grid, nbblk;
  int bulk_ok, bulk_grid;   // every chunk fits a bulk stage; persistent grid of the bulk-copy kernels
};
namespace pyn {
static inline SpmvPlanView view_of(const pyn_spmv_plan* p) {
  SpmvPlanView v;
  v.nrows = p->nrows; v.nnz = p->nnz; v.ptr = p->ptr; v.blk_row = p->blk_row; v.nblk = p->nblk;
  v.brow = p->brow; v.bp0 = p->bp0; v.nbblk = p->nbblk;
  return v;
}
}  // namespace pyn
