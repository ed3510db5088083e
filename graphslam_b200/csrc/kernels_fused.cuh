// Small-graph PCG: the whole iteration in ONE persistent kernel (single GPU).
//
// On graphs that fit the chip in one wave (every tile's CTA resident at once: up to ~227k poses) a PCG
// iteration of three launches is launch-latency bound: ~17 us per iteration on the 10k-50k-pose graphs of
// the incremental mode (C5), of which the kernels' own work is a few microseconds.  Here one cooperative
// launch runs a whole batch of iterations; the phases are separated by a grid barrier instead of a kernel
// boundary, and after each barrier EVERY CTA folds the tile partials itself (same fixed tree:
// tile -> segment -> total) so no second barrier is needed to publish alpha / beta.
//
// Tried on top of this and dropped (B200, C5 12k / 36k poses, C2 100k; seconds per 3 LM steps):
//   * taking a row's slots 2 or 4 at a time (more loads in flight per row): 0.129-0.133 against 0.133 / 0.244-
//     0.253 against 0.254 — the phases are bounded by the barrier, not by the row's chain of loads;
//   * making the barrier the reduction (every CTA publishes its partial as flagged words and polls everybody
//     else's instead of arriving on a counter): 0.165 / 0.387 / C2 1.05 against 0.133 / 0.254 / 0.61 — N^2 polls.
//
// The arithmetic is the very same per-row code (rows.cuh) and the same reduction tree as the three-kernel
// path, so the iterates are bit-identical to it (tests/test_gpu_parity.py::test_fused_pcg_bitwise).
#pragma once
#include "kernels.cuh"

namespace gs {

struct GridBarrier {
  unsigned int count;
  unsigned int gen;
};

__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Sense-reversing central barrier; all CTAs are co-resident (cooperative launch).  The trailing gpu-scope
// fence also drops the SM's L1 lines, so plain loads after the barrier see what other CTAs wrote before it.
__device__ __forceinline__ void grid_sync(GridBarrier* b) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int gen = ld_acquire_gpu(&b->gen);
    __threadfence();
    if (atomicAdd(&b->count, 1u) == gridDim.x - 1) {
      b->count = 0u;
      __threadfence();
      atomicAdd(&b->gen, 1u);
    } else {
      while (ld_acquire_gpu(&b->gen) == gen) {
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// every CTA: tile partials -> 8 segment sums (warp w folds segment w exactly as reduce_finish's last CTA
// does) -> totals in segment order, left in tot[] for the whole CTA
template <int NQ>
__device__ __forceinline__ void fold_all(const double* part, int part_stride, const int* seg_tile, double (*segs)[kSegments],
                                         double* tot) {
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (w < kSegments) {
    const int t0 = seg_tile[w], t1 = seg_tile[w + 1];
    double acc[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) acc[q] = 0.0;
    for (int i = t0 + l; i < t1; i += 32) {
#pragma unroll
      for (int q = 0; q < NQ; ++q) acc[q] += __ldcg(part + (size_t)q * part_stride + i);
    }
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      const double t = warp_sum(acc[q]);
      if (l == 0) segs[q][w] = t;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      double t = 0.0;
      for (int g = 0; g < kSegments; ++g) t += segs[q][g];
      tot[q] = t;
    }
  }
  __syncthreads();
}

// part: 3 rows of part_stride doubles (rows 0,1: r.z and r.r, row 2: p.Ap — separate rows, so a CTA that is
// already writing the next phase's partials cannot overwrite what a slower CTA is still folding).
//
// Two barriers per iteration, not three: the direction update p <- z + beta p is not a phase of its own.  The
// product forms p(c) = z(c) + beta pold(c) on the fly for every column it touches and each row stores its own new
// p into the OTHER of two p buffers, so nobody reads a p that somebody else is writing.  An even batch leaves the
// latest direction in the first buffer (= the graph's p vector) for the next launch; beta travels in Scalars
// (0 at the start of a solve, where p = z already).
__global__ void __launch_bounds__(kThreads, 4) k_pcg_fused(const SpmvView sv, const PcgView a, double* Ap, double* p_alt,
                                                         double* __restrict__ part, int part_stride,
                                                         const int* __restrict__ seg_tile, Scalars* sc, GridBarrier* bar,
                                                         int batch) {
  __shared__ double red[8];
  __shared__ double segs[2][kSegments];
  __shared__ double tot[2];
  const int r = blockIdx.x * kThreads + threadIdx.x;
  const bool live = r < a.n_own;
  // the scalar state, identical in every CTA
  int done = sc->done;
  if (done) return;   // uniform
  double rz = sc->rz;
  const double rr0 = sc->rr0, tol2 = sc->tol2, lambda = sc->lambda;
  const int max_iters = sc->max_iters;
  int iters = sc->iters;
  double pAp = sc->pAp, rr = sc->rr, beta = sc->beta;
  double* p_prev = a.p;    // holds the previous direction
  double* p_next = p_alt;  // receives this iteration's
#pragma unroll 1
  for (int k = 0; k < batch; ++k) {
    // ---- A: p = z + beta p_prev (formed on the fly, own row stored to the other buffer), Ap = (H + lambda D) p, p.Ap
    {
      const double v = live ? spmv_row_t<true>(sv, p_prev, nullptr, a.z, nullptr, beta, p_next, Ap, r, lambda) : 0.0;
      const double s = block_sum(v, red);
      if (threadIdx.x == 0) part[2 * (size_t)part_stride + blockIdx.x] = s;
    }
    {
      double* t = p_prev;
      p_prev = p_next;   // from here on p_prev is the current direction
      p_next = t;
    }
    grid_sync(bar);
    fold_all<1>(part + 2 * (size_t)part_stride, part_stride, seg_tile, segs, tot);
    pAp = tot[0];
    if (!(pAp > 0.0)) {
      done = 2;
      break;
    }
    const double alpha = rz / pAp;
    // ---- B: x += alpha p ; r -= alpha Ap ; z = M^-1 r ; r.z, r.r
    {
      PcgView b = a;
      b.p = p_prev;
      b.Ap = Ap;
      double v0 = 0.0, v1 = 0.0;
      if (live) pcg_update_row(b, r, alpha, v0, v1);
      const double s0 = block_sum(v0, red);
      if (threadIdx.x == 0) part[blockIdx.x] = s0;
      const double s1 = block_sum(v1, red);
      if (threadIdx.x == 0) part[(size_t)part_stride + blockIdx.x] = s1;
    }
    grid_sync(bar);
    fold_all<2>(part, part_stride, seg_tile, segs, tot);
    beta = tot[0] / rz;
    rz = tot[0];
    rr = tot[1];
    iters += 1;
    if (rr <= tol2 * rr0) done = 1;
    else if (!(rr == rr)) done = 2;
    else if (iters >= max_iters) done = 3;
    if (done) break;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    sc->pAp = pAp;
    sc->rz = rz;
    sc->rr = rr;
    sc->beta = beta;
    sc->iters = iters;
    sc->done = done;
  }
}

}  // namespace gs
