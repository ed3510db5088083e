// sm_100a kernels of the Gauss-Newton / Levenberg-Marquardt iteration (FP64, HBM-bound, no tensor
// cores: no stage is a dense contraction).  One thread per block row; the per-row bodies are in
// rows.cuh.  Every reduction runs tile (one CTA, fixed shuffle tree) -> segment (fixed-order sum of
// the tile partials by the last CTA to finish) -> total (the kSegments segment sums added in
// order), so results are bit-reproducible run to run AND identical for 1, 2, 4 or 8 ranks.
//
//   K1+K2  k_assemble        per-factor SE(2) linearisation (BetweenFactor / PriorFactor + Gaussian
//                            whitening) fused with the row-owner reduction into H = sum J^T Omega J
//                            (diagonal planes + one stored block per vertex pair) and b
//   K3     k_spmv            y = (H + lambda*D) p over the symmetric block storage, fused p^T y
//   K4     k_pcg_init / k_pcg_update / k_pcg_dir   block-Jacobi PCG vector updates with fused dots
//   K5     k_retract, k_chi2 Values::retract and graph.error
//
// What each replaces in the reference is GTSAM code reached from
// LevenbergMarquardtOptimizer::optimize() (reference src/graph/src/graph.cpp:119); see DESIGN.md.
#pragma once
#include <cstdint>
#include "rows.cuh"

namespace gs {

constexpr int kThreads = kTile;

// Minimum resident CTAs per SM of the product kernels = their register budget.  Left to itself ptxas compiles the
// slot loop to 40 registers and gets there by reusing the registers of the first loaded entries for the next ones:
// the twelve loads of a slot then go out in FOUR dependent groups (profiles/r2_spmv_sass.md), four memory round
// trips per slot in a kernel that is latency-bound (75 % long-scoreboard stalls at 48 % DRAM utilisation).  With
// 48 registers (five CTAs) all twelve are requested before the first use.
#ifndef GS_SPMV_MINB
#define GS_SPMV_MINB 4
#endif
#ifndef GS_SPMVDIR_MINB
#define GS_SPMVDIR_MINB 4
#endif

// Device-resident scalars: PCG coefficients never visit the host.
struct Scalars {
  double rz;        // r^T z of the current iterate
  double pAp;
  double rr;        // |r|^2
  double rr0;       // |g|^2
  double beta;
  double lambda;
  double tol2;      // pcg_rel_tol^2
  double chi2;      // k_chi2 / k_assemble result
  double dot;       // k_dot / plain k_spmv result
  int done;         // 0 running, 1 converged, 2 breakdown (p^T A p <= 0), 3 max iters, 4 peer wait timed out
  int iters;
  int max_iters;
  unsigned int seq0;  // peer-memory path: sequence number of this solve's first iteration
  double rzv[2];      // peer-memory path: r^T z by iteration parity (reader and writer never share a slot)
  long long spin[4];  // peer-memory path: clock ticks CTA 0 spent waiting in spmv / update / dir, and the waits counted
  double rz_bj;       // two-level preconditioner: r.M^-1 r of the new residual, before the coarse term c.mu is added
};

enum ReduceOp { OP_PAP = 0, OP_DOT = 1, OP_INIT = 2, OP_UPDATE = 3, OP_CHI2 = 4, OP_INIT_2L = 5, OP_UPDATE_2L = 6 };

// where a kernel's reduction goes
struct ReduceCtx {
  double* part;            // [nq][part_stride] per-tile partials
  int part_stride;
  double* segs;            // [world][nq][spr] segment sums; this rank fills its own chunk
  const int* seg_tile;     // spr+1 local tile boundaries
  int spr;                 // segments per rank
  int rank;
  int world;
  unsigned int* counter;
  Scalars* sc;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Fixed-tree block sum; result valid in warp 0.
__device__ __forceinline__ double block_sum(double v, double* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  double t = 0.0;
  if (w == 0) {
    t = (l < (int)(blockDim.x >> 5)) ? red[l] : 0.0;
    t = warp_sum(t);
  }
  return t;
}

// True in every thread of the CTA that finishes last (its partial is already published).
__device__ __forceinline__ bool last_block(unsigned int* counter, int* flag) {
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(counter, 1u);
    *flag = (t == gridDim.x - 1);
  }
  __syncthreads();
  return *flag != 0;
}

// Fixed-order sum of n per-CTA partials by one CTA (thread t takes t, t+T, ...; then fixed tree).
__device__ __forceinline__ double final_sum(const double* part, int n, double* red) {
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += __ldcg(part + i);
  return block_sum(s, red);
}

// total of quantity q = the kSegments segment sums in global segment order
__device__ __forceinline__ double seg_total(const double* segs, int q, int nq, int spr) {
  double t = 0.0;
  for (int g = 0; g < kSegments; ++g) t += __ldcg(segs + ((size_t)(g / spr) * nq + q) * spr + (g % spr));
  return t;
}

// the stateful scalar step that follows each reduction (one thread)
__device__ __forceinline__ void apply_scalars(Scalars* sc, int op, double t0, double t1) {
  switch (op) {
    case OP_PAP:
      if (sc->done) return;
      sc->pAp = t0;
      if (!(t0 > 0.0)) sc->done = 2;
      break;
    case OP_DOT:
      sc->dot = t0;
      break;
    case OP_INIT:
      sc->rz = t0;
      sc->rzv[0] = t0;
      sc->rzv[1] = t0;
      sc->rr = t1;
      sc->rr0 = t1;
      sc->beta = 0.0;
      sc->iters = 0;
      sc->done = (t1 == 0.0) ? 1 : 0;
      break;
    case OP_UPDATE: {
      if (sc->done) return;
      sc->beta = t0 / sc->rz;
      sc->rz = t0;
      sc->rr = t1;
      const int it = sc->iters + 1;
      sc->iters = it;
      if (t1 <= sc->tol2 * sc->rr0) sc->done = 1;
      else if (!(t1 == t1)) sc->done = 2;
      else if (it >= sc->max_iters) sc->done = 3;
      break;
    }
    case OP_CHI2:
      sc->chi2 = t0;
      break;
    // two-level preconditioner: r.z is completed by k_coarse_mv (r.M^-1 r here, + c.mu there), which also forms beta
    case OP_INIT_2L:
      sc->rz_bj = t0;
      sc->rr = t1;
      sc->rr0 = t1;
      sc->beta = 0.0;
      sc->iters = 0;
      sc->done = (t1 == 0.0) ? 1 : 0;
      break;
    case OP_UPDATE_2L: {
      if (sc->done) return;
      sc->rz_bj = t0;
      sc->rr = t1;
      const int it = sc->iters + 1;
      sc->iters = it;
      if (t1 <= sc->tol2 * sc->rr0) sc->done = 1;
      else if (!(t1 == t1)) sc->done = 2;
      else if (it >= sc->max_iters) sc->done = 3;
      break;
    }
  }
}

// tile partials -> (last CTA) segment sums -> (single rank) totals and scalar step.
// In the last CTA warp w folds segment w: lane l adds partials l, l+32, ... in order, then the fixed
// butterfly; all NQ quantities in one pass, no block barrier per segment.
template <int NQ>
__device__ __forceinline__ void reduce_finish(const ReduceCtx& c, const double (&v)[NQ], int op, double* red, int* flag) {
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const double s = block_sum(v[q], red);
    if (threadIdx.x == 0) c.part[(size_t)q * c.part_stride + blockIdx.x] = s;
  }
  if (last_block(c.counter, flag)) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    for (int s = w; s < c.spr; s += (int)(blockDim.x >> 5)) {
      const int t0 = c.seg_tile[s], t1 = c.seg_tile[s + 1];
      double acc[NQ];
#pragma unroll
      for (int q = 0; q < NQ; ++q) acc[q] = 0.0;
      for (int i = t0 + l; i < t1; i += 32) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) acc[q] += __ldcg(c.part + (size_t)q * c.part_stride + i);
      }
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const double t = warp_sum(acc[q]);
        if (l == 0) c.segs[((size_t)c.rank * NQ + q) * c.spr + s] = t;
      }
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      if (c.world == 1) {
        const double a = seg_total(c.segs, 0, NQ, c.spr);
        const double b = NQ > 1 ? seg_total(c.segs, 1, NQ, c.spr) : 0.0;
        apply_scalars(c.sc, op, a, b);
      }
      *c.counter = 0u;
    }
  }
}

// multi-rank: runs after the all-gather of the segment sums
__global__ void k_scalars(const double* segs, int nq, int spr, int op, Scalars* sc) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const double a = seg_total(segs, 0, nq, spr);
    const double b = nq > 1 ? seg_total(segs, 1, nq, spr) : 0.0;
    apply_scalars(sc, op, a, b);
  }
}

// ------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(kThreads, 2) k_assemble(const AsmView a, const ReduceCtx c) {
  __shared__ double red[8];
  __shared__ int flag;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < a.n_own) v[0] = assemble_row<MODE>(a, r);
  reduce_finish<1>(c, v, OP_CHI2, red, &flag);
}

// Candidate for round 2 (GS_ASM_NODUP=1, off by default, not yet timed on a B200): graphs without duplicate
// factors need no run accumulator, and without it the row fits three CTAs per SM instead of two (k_assemble is
// bound by its 4 warps per scheduler).  Same bits as k_assemble on such graphs
// (tests/test_emulator_cpu.py::test_assemble_without_run_accumulator_is_bitwise_the_same).
template <int MODE>
__global__ void __launch_bounds__(kThreads, 3) k_assemble_nodup(const AsmView a, const ReduceCtx c) {
  __shared__ double red[8];
  __shared__ int flag;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < a.n_own) v[0] = assemble_row<MODE, true>(a, r);
  reduce_finish<1>(c, v, OP_CHI2, red, &flag);
}

template <int PIPE>
__global__ void __launch_bounds__(kThreads) k_chi2(const AsmView a, const ReduceCtx c) {
  __shared__ double red[8];
  __shared__ int flag;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < a.n_own) v[0] = chi2_row<PIPE>(a, r);
  reduce_finish<1>(c, v, OP_CHI2, red, &flag);
}

// mode 0: PCG step (lambda from sc, honours sc->done, result -> sc->pAp)
// mode 1: plain undamped product (result -> sc->dot)
// (A variant that staged several slots per thread through shared memory with cp.async was measured
//  slower on C3 — 123-165 us against 101 us — because the staging buffers cost occupancy and L1; see
//  DESIGN.md "What did not help".)
__global__ void __launch_bounds__(kThreads, GS_SPMV_MINB) k_spmv(const SpmvView a, const double* __restrict__ p, const double* __restrict__ pg,
                                                    double* __restrict__ y, const ReduceCtx c, int mode) {
  __shared__ double red[8];
  __shared__ int flag;
  if (mode == 0 && c.sc->done) return;
  const double lambda = (mode == 0) ? c.sc->lambda : 0.0;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < a.n_own) v[0] = spmv_row(a, p, pg, y, r, lambda);
  reduce_finish<1>(c, v, mode == 0 ? OP_PAP : OP_DOT, red, &flag);
}

__global__ void __launch_bounds__(kThreads) k_minv(const double* __restrict__ Hd, const double* __restrict__ dscale,
                                                    const Scalars* sc, double* __restrict__ Minv, int n_own) {
  const int v = blockIdx.x * kThreads + threadIdx.x;
  if (v < n_own) minv_row(Hd, dscale, Minv, n_own, v, sc->lambda);
}

// diagonalDamping: clip(diag(H), 1e-6, 1e32)  (GTSAM min_diagonal / max_diagonal)
__global__ void __launch_bounds__(kThreads) k_dscale(const double* __restrict__ Hd, double* __restrict__ dscale, int n_own) {
  const int v = blockIdx.x * kThreads + threadIdx.x;
  if (v >= n_own) return;
  const size_t n = (size_t)n_own;
  dscale[3 * (size_t)v] = fmin(fmax(Hd[v], 1e-6), 1e32);
  dscale[3 * (size_t)v + 1] = fmin(fmax(Hd[3 * n + v], 1e-6), 1e32);
  dscale[3 * (size_t)v + 2] = fmin(fmax(Hd[5 * n + v], 1e-6), 1e32);
}

__global__ void __launch_bounds__(kThreads) k_pcg_init(const PcgView a, const ReduceCtx c, int op) {
  __shared__ double red[8];
  __shared__ int flag;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[2] = {0.0, 0.0};
  if (r < a.n_own) pcg_init_row(a, r, v[0], v[1]);
  reduce_finish<2>(c, v, op, red, &flag);
}

// The x/r/z update with every global access a full-width one.
// pcg_update_row reads and writes its vectors three doubles per thread at a 24-byte stride: a warp's store
// instruction covers a third of each 32-byte sector it touches, so the nine stores of a row reach L2 as three
// partial writes per sector, and the twelve loads as three requests per line.  Here the tile's 768 doubles of
// x, p, r, Ap move as double2 (one or two per thread, consecutive threads on consecutive 16 bytes); x and r are
// elementwise, the new r goes through shared memory to the thread that owns the row for z = M^-1 r and the two
// dot products, and z goes back through shared memory to be stored flat.  Same expressions per element and per
// row as pcg_update_apply, same reduction tree: same bits (tests/test_gpu_parity.py::test_staged_update_bitwise).
// C3: 51.5 -> 47.6 us per launch (profiles/r1_v7_variant_probe.jsonl).
__global__ void __launch_bounds__(kThreads) k_pcg_update_staged(const PcgView a, const ReduceCtx c, int op) {
  __shared__ double red[8];
  __shared__ int flag;
  __shared__ __align__(16) double sv[3 * kThreads];
  if (c.sc->done) return;
  const double alpha = c.sc->rz / c.sc->pAp;
  const int row0 = blockIdx.x * kThreads;
  const int nrow = min(kThreads, a.n_own - row0);        // rows of this tile (the last one may be short)
  const size_t base = 3 * (size_t)row0;                  // 768 doubles per tile: 16-byte aligned
  const int nd = 3 * nrow, nd2 = nd >> 1;
  const double2* p2 = reinterpret_cast<const double2*>(a.p + base);
  const double2* q2 = reinterpret_cast<const double2*>(a.Ap + base);
  double2* x2 = reinterpret_cast<double2*>(a.x + base);
  double2* r2 = reinterpret_cast<double2*>(a.r + base);
  double2* s2 = reinterpret_cast<double2*>(sv);
  for (int i = threadIdx.x; i < nd2; i += kThreads) {   // at most two trips
    const double2 pv = p2[i], qv = q2[i], xv = x2[i], rv = r2[i];
    const double2 xn = make_double2(xv.x + alpha * pv.x, xv.y + alpha * pv.y);
    const double2 rn = make_double2(rv.x - alpha * qv.x, rv.y - alpha * qv.y);
    x2[i] = xn;
    r2[i] = rn;
    s2[i] = rn;
  }
  if ((nd & 1) && threadIdx.x == 0) {                    // odd number of rows in a short last tile
    const size_t e = base + (size_t)nd - 1;
    a.x[e] = a.x[e] + alpha * a.p[e];
    const double rn = a.r[e] - alpha * a.Ap[e];
    a.r[e] = rn;
    sv[nd - 1] = rn;
  }
  __syncthreads();
  double v[2] = {0.0, 0.0};
  double z0 = 0.0, z1 = 0.0, z2 = 0.0;
  const int t = threadIdx.x;
  if (t < nrow) {
    const double r0 = sv[3 * t], r1 = sv[3 * t + 1], r2v = sv[3 * t + 2];
    // (requesting the row's M^-1 block before the barrier, together with the tile's vectors, costs 18 registers and
    //  two resident CTAs: 71 us against 48 us on C3; profiles/r2_variant_probe.md)
    apply_minv(a.Minv, (size_t)a.n_own, (size_t)(row0 + t), r0, r1, r2v, z0, z1, z2);
    v[0] = (r0 * z0 + r1 * z1) + r2v * z2;
    v[1] = (r0 * r0 + r1 * r1) + r2v * r2v;
  }
  __syncthreads();                                       // every row has read its r
  if (t < nrow) {
    sv[3 * t] = z0;
    sv[3 * t + 1] = z1;
    sv[3 * t + 2] = z2;
  }
  __syncthreads();
  double2* z2p = reinterpret_cast<double2*>(a.z + base);
  for (int i = threadIdx.x; i < nd2; i += kThreads) z2p[i] = s2[i];
  if ((nd & 1) && threadIdx.x == 0) a.z[base + (size_t)nd - 1] = sv[nd - 1];
  reduce_finish<2>(c, v, op, red, &flag);
}

// p = z + beta p over the flat owned range: four doubles per thread as two 16-byte accesses (one 8-byte
// access per thread left the kernel at 0.66 of the HBM peak); n3 = 3 n_own, the vectors are 256-byte aligned
constexpr int kDirPerThread = 4;
__global__ void __launch_bounds__(kThreads) k_pcg_dir(const double* __restrict__ z, double* __restrict__ p,
                                                       const Scalars* sc, int n3) {
  if (sc->done) return;
  const double beta = sc->beta;
  const int i = (blockIdx.x * kThreads + threadIdx.x) * 2;                       // first double2 of this thread
  const int j = i + (int)gridDim.x * kThreads * 2;                               // second one, a grid away
  const int n2 = n3 & ~1;
  double2 za = make_double2(0.0, 0.0), pa = za, zb = za, pb = za;
  if (i < n2) {
    za = *reinterpret_cast<const double2*>(z + i);
    pa = *reinterpret_cast<const double2*>(p + i);
  }
  if (j < n2) {
    zb = *reinterpret_cast<const double2*>(z + j);
    pb = *reinterpret_cast<const double2*>(p + j);
  }
  if (i < n2)
    *reinterpret_cast<double2*>(p + i) = make_double2(dir_value(za.x, beta, pa.x), dir_value(za.y, beta, pa.y));
  if (j < n2)
    *reinterpret_cast<double2*>(p + j) = make_double2(dir_value(zb.x, beta, pb.x), dir_value(zb.y, beta, pb.y));
  if ((n3 & 1) && blockIdx.x == 0 && threadIdx.x == 0) p[n3 - 1] = dir_value(z[n3 - 1], beta, p[n3 - 1]);
}

__global__ void __launch_bounds__(kThreads) k_dot(const double* __restrict__ u, const double* __restrict__ w, int n_own,
                                                   const ReduceCtx c) {
  __shared__ double red[8];
  __shared__ int flag;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < n_own) v[0] = dot_row(u, w, r);
  reduce_finish<1>(c, v, OP_DOT, red, &flag);
}

__global__ void __launch_bounds__(kThreads) k_retract(const double4* __restrict__ pose, const double* __restrict__ d,
                                                       double4* __restrict__ out, int n_own) {
  const int v = blockIdx.x * kThreads + threadIdx.x;
  if (v >= n_own) return;
  const size_t o = 3 * (size_t)v;
  out[v] = retract(pose[v], d[o], d[o + 1], d[o + 2]);
}

// halo pack: out[k*width + c] = vec[rows[k]*width + c]   (width 3: PCG vectors, 4: poses)
__global__ void __launch_bounds__(kThreads) k_pack(const double* __restrict__ vec, const int* __restrict__ rows, int n,
                                                    int width, double* __restrict__ out) {
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= n * width) return;
  const int k = i / width, c = i - k * width;
  out[i] = vec[(size_t)rows[k] * width + c];
}

// closest_keyframe (reference graph.cpp:146-178): argmin of the Euclidean distance to the query pose over the first
// `count` vertices in insertion order; ties -> lowest insertion index (the reference's strict '<', graph.cpp:162).
// Grid-wide: every CTA scans a tile of candidates (32 B per vertex, coalesced when keys were inserted in order),
// reduces (distance, index) lexicographically, and the last CTA to finish folds the per-CTA results.  The query pose
// is read on the device (no host round trip).  On several ranks each rank scans the candidates it owns (dense rows
// [v0, v0 + n_own)); the host then takes the minimum of the ranks' results with the same tie rule.
struct ClosestOut {
  double dist;
  int idx;      // insertion index, 0x7fffffff = no candidate on this rank
  int pad;
};

constexpr int kClosestPerThread = 8;

__device__ __forceinline__ void closest_min(double& bd, int& bi, double od, int oi) {
  if (od < bd || (od == bd && oi < bi)) {
    bd = od;
    bi = oi;
  }
}

__global__ void __launch_bounds__(kThreads) k_closest(const double4* __restrict__ pose, const int* __restrict__ order, int count,
                                                       const double4* __restrict__ query, int v0, int n_own,
                                                       ClosestOut* __restrict__ part, unsigned int* counter,
                                                       ClosestOut* __restrict__ out) {
  __shared__ double sd[kThreads / 32];
  __shared__ int si[kThreads / 32];
  __shared__ int flag;
  const double4 q = ldg(query);
  double bd = INFINITY;
  int bi = 0x7fffffff;
  const int base = blockIdx.x * (kThreads * kClosestPerThread) + threadIdx.x;
#pragma unroll
  for (int k = 0; k < kClosestPerThread; ++k) {
    const int t = base + k * kThreads;
    if (t < count) {
      const int dense = order ? order[t] : t;
      const int local = dense - v0;
      if (local >= 0 && local < n_own) {
        const double4 p = ldg(pose + local);
        // sqrt(pow(dx,2)+pow(dy,2)) as the reference computes it (graph.cpp:157)
        const double dist = sqrt((p.x - q.x) * (p.x - q.x) + (p.y - q.y) * (p.y - q.y));
        closest_min(bd, bi, dist, t);
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) closest_min(bd, bi, __shfl_xor_sync(0xffffffffu, bd, o), __shfl_xor_sync(0xffffffffu, bi, o));
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) {
    sd[w] = bd;
    si[w] = bi;
  }
  __syncthreads();
  if (w == 0) {
    bd = l < kThreads / 32 ? sd[l] : INFINITY;
    bi = l < kThreads / 32 ? si[l] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) closest_min(bd, bi, __shfl_xor_sync(0xffffffffu, bd, o), __shfl_xor_sync(0xffffffffu, bi, o));
    if (l == 0) {
      part[blockIdx.x].dist = bd;
      part[blockIdx.x].idx = bi;
    }
  }
  if (!last_block(counter, &flag)) return;
  bd = INFINITY;
  bi = 0x7fffffff;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += kThreads) {
    const double od = __ldcg(&part[i].dist);
    const int oi = __ldcg(&part[i].idx);
    closest_min(bd, bi, od, oi);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) closest_min(bd, bi, __shfl_xor_sync(0xffffffffu, bd, o), __shfl_xor_sync(0xffffffffu, bi, o));
  __syncthreads();
  if (l == 0) {
    sd[w] = bd;
    si[w] = bi;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < kThreads / 32; ++k) closest_min(bd, bi, sd[k], si[k]);
    out->dist = bd;
    out->idx = bi;
    *counter = 0u;
  }
}

}  // namespace gs
