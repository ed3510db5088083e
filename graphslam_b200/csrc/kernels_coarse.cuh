// Coarse-space correction of the block-Jacobi preconditioner (two-level additive preconditioner):
//
//     z = M^-1 r  +  W E^-1 W^T r ,      E = W^T (H + lambda I) W
//
// W has three columns per AGGREGATE of about a thousand graph-adjacent poses (structure.hpp: build_aggregation): the
// rigid-body motions of the aggregate — world-frame translation (2) and rotation about the aggregate's reference
// pose (1) — written in the body-frame tangent coordinates GTSAM's Pose2 retract uses:
//
//     W_i = [ R_i^T   R_i^T J (t_i - t_ref) ]        J = [0 -1; 1 0]
//           [ 0  0            1             ]
//
// They span the near-null space of the between-factors' Hessian (a rigid motion of a whole aggregate changes no
// residual inside it), i.e. exactly the smooth error block-Jacobi PCG needs thousands of iterations for: on the
// 1M-pose graph 14,700 iterations per solve become about 700.  W is never stored: a row's
// block follows from its pose and the reference pose.  Everything is deterministic (fixed-order sums, no atomics).
//
//   k_coarse_product  Y = H W_c for the three columns of W of one colour's aggregates (set-up: E's columns), k_restrict_modes
//   k_basis_vec   v = column `mode` of W (set-up: G = W^T W)
//   k_restrict    c_a = sum_{i in a} W_i^T r_i                       one CTA per aggregate, rows in list order
//   k_coarse_E    dense E = E0 + lambda G from the restricted products (layout: column-major, 3 n_agg square)
//   k_coarse_mv   mu = E^-1 c (dense symmetric product, one warp per row) and c.mu, then the scalar step
//   k_dir_coarse  p = (z + W mu) + beta p                            the direction update with the prolongation folded in
#pragma once
#include "kernels.cuh"

namespace gs {

// Aggregates never cross segment boundaries, so a rank owns the contiguous range [a0, a0 + n_agg_own) of the global
// aggregate numbering; coarse vectors (3 per aggregate) are indexed globally, lists and rows locally.
struct CoarseView {
  const int* agg_of;        // per local row (owned rows, then ghost rows): GLOBAL aggregate
  const int* agg_ptr;       // per OWNED aggregate (local index): range in agg_rows
  const int* agg_rows;      // local rows, ascending; the first one is the aggregate's reference row
  const double4* pose;      // the linearisation point
  const double2* ref;       // per global aggregate: position of its reference row (several ranks: gathered per linearisation)
  int a0;                   // first owned aggregate
  int n_agg_own;
  int n_agg;                // global
  int n_own;
};

// W_i as four numbers: (c, s) of the pose, (u, v) = R^T J (t_i - t_ref)
struct WBlock {
  double c, s, u, v;
};

__device__ __forceinline__ WBlock w_block(const double4 p, const double2 ref) {
  const double dx = p.x - ref.x, dy = p.y - ref.y;
  WBlock w;
  w.c = p.z;
  w.s = p.w;
  w.u = p.w * dx - p.z * dy;    // R^T (-dy, dx), first component
  w.v = p.w * dy + p.z * dx;
  return w;
}

// column `mode` of W_i
__device__ __forceinline__ void w_column(const WBlock& w, int mode, double out[3]) {
  if (mode == 0) { out[0] = w.c; out[1] = -w.s; out[2] = 0.0; }
  else if (mode == 1) { out[0] = w.s; out[1] = w.c; out[2] = 0.0; }
  else { out[0] = w.u; out[1] = w.v; out[2] = 1.0; }
}

// reference position of every aggregate = its first row's
__global__ void __launch_bounds__(kThreads) k_agg_ref(const CoarseView a, double2* __restrict__ ref) {
  const int g = blockIdx.x * kThreads + threadIdx.x;
  if (g >= a.n_agg_own) return;
  const double4 p = ldg(a.pose + a.agg_rows[a.agg_ptr[g]]);
  ref[a.a0 + g] = make_double2(p.x, p.y);
}

// color < 0: every aggregate.  rows = n_own, or n_own + n_ghost when the product that follows reads the ghost rows
__global__ void __launch_bounds__(kThreads) k_basis_vec(const CoarseView a, int rows, const int* __restrict__ agg_color, int color,
                                                         int mode, double* __restrict__ v) {
  const int r = blockIdx.x * kThreads + threadIdx.x;
  if (r >= rows) return;
  const int g = a.agg_of[r];
  double o[3] = {0.0, 0.0, 0.0};
  if (color < 0 || agg_color[g] == color) {
    w_column(w_block(ldg(a.pose + r), ldg(a.ref + g)), mode, o);
  }
  v[3 * (size_t)r] = o[0];
  v[3 * (size_t)r + 1] = o[1];
  v[3 * (size_t)r + 2] = o[2];
}

// this thread's share of sum_{i in aggregate g} W_i^T r_i (g = local aggregate index): rows threadIdx.x, + T, ... of
// the aggregate's list, in list order — ONE spelling for the single-GPU and the peer-memory kernel
// (STRIDE doubles between consecutive rows of r: 3 for the PCG vectors, 9 for the three-column products of the set-up)
template <int STRIDE = 3>
__device__ __forceinline__ void restrict_sums(const CoarseView& a, const double* __restrict__ r, int g, double& s0, double& s1,
                                              double& s2) {
  const int q0 = a.agg_ptr[g], q1 = a.agg_ptr[g + 1];
  const double2 ref = ldg(a.ref + a.a0 + g);
  s0 = 0.0;
  s1 = 0.0;
  s2 = 0.0;
  // four rows per thread and trip: the four row indices, then their poses and residuals, are requested together;
  // the sums run over the thread's rows in list order
  for (int q = q0 + (int)threadIdx.x; q < q1; q += 4 * kThreads) {
    int idx[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) idx[u] = (q + u * kThreads < q1) ? ldg(a.agg_rows + q + u * kThreads) : -1;
    double4 ps[4];
    double rv[4][3];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = idx[u] < 0 ? 0 : idx[u];
      ps[u] = ldg(a.pose + i);
      rv[u][0] = r[STRIDE * (size_t)i];
      rv[u][1] = r[STRIDE * (size_t)i + 1];
      rv[u][2] = r[STRIDE * (size_t)i + 2];
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (idx[u] < 0) continue;
      const WBlock w = w_block(ps[u], ref);
      s0 += w.c * rv[u][0] - w.s * rv[u][1];
      s1 += w.s * rv[u][0] + w.c * rv[u][1];
      s2 += (w.u * rv[u][0] + w.v * rv[u][1]) + rv[u][2];
    }
  }
}

// out[a * ostride + ooff .. + 3) = sum over the rows of owned aggregate a (list order) of W_i^T r_i (a = global index).
// Thread t takes rows t, t+T, ...; then the fixed shuffle / shared-memory tree of block_sum.
__global__ void __launch_bounds__(kThreads, 4) k_restrict(const CoarseView a, const double* __restrict__ r, const Scalars* sc,
                                                        int honour_done, double* __restrict__ out, int ostride, int ooff) {
  __shared__ double red[8];
  if (honour_done && sc->done) return;
  const int g = blockIdx.x;                   // local aggregate
  double s0, s1, s2;
  restrict_sums(a, r, g, s0, s1, s2);
  const double t0 = block_sum(s0, red);
  const double t1 = block_sum(s1, red);
  const double t2 = block_sum(s2, red);
  if (threadIdx.x == 0) {
    double* o = out + (size_t)(a.a0 + g) * ostride + ooff;
    o[0] = t0;
    o[1] = t1;
    o[2] = t2;
  }
}

// Set-up: the three columns of E0 = W^T H W that belong to the aggregates of ONE colour, in one pass over H.
// Y[9 r + 3 mode + k] = (H w_mode)_r,k where w_mode is column `mode` of W restricted to the aggregates of that colour: a
// column's block of W is formed on the fly from its pose and its aggregate's reference (ghost columns included: the
// rank holds their poses), so there is no basis vector in memory, no exchange, and H is read once instead of three
// times.  Per right-hand side the arithmetic is the product kernel's (spmv_slot per stored block in slot order, then the
// diagonal block).
__global__ void __launch_bounds__(kThreads) k_coarse_product(const SpmvView a, const CoarseView cv, const int* __restrict__ agg_color,
                                                              int color, double* __restrict__ Y) {
  const int r = blockIdx.x * kThreads + threadIdx.x;
  if (r >= a.n_own) return;
  const int sl = r / kSlice, lane = r % kSlice;
  const int base = a.rs.base[sl] + lane, w = a.rs.width[sl];
  const size_t n = (size_t)a.n_own, rr = (size_t)r;
  double y[3][3] = {{0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}, {0.0, 0.0, 0.0}};
  double d[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) d[k] = ldg(a.Hd + (size_t)k * n + rr);
  int2 nref = w > 0 ? ldg(a.refs + base) : make_int2(-1, 0);
  for (int s = 0; s < w; ++s) {
    const int2 ref = nref;
    if (s + 1 < w) nref = ldg(a.refs + base + (s + 1) * kSlice);
    if (ref.x < 0) continue;
    const int g = cv.agg_of[ref.x];
    if (agg_color[g] != color) continue;          // a zero column adds nothing
    double h[9];
    load_block_as(a.Ho, a.hstride, ref.y >> 1, ref.y & 1, h);
    const WBlock wb = w_block(ldg(cv.pose + ref.x), ldg(cv.ref + g));
#pragma unroll
    for (int mode = 0; mode < 3; ++mode) {
      double v[3];
      w_column(wb, mode, v);
      spmv_slot(h, v[0], v[1], v[2], y[mode][0], y[mode][1], y[mode][2]);
    }
  }
  const int g = cv.agg_of[r];
  const bool mine = agg_color[g] == color;
  const WBlock wb = w_block(ldg(cv.pose + r), ldg(cv.ref + g));
#pragma unroll
  for (int mode = 0; mode < 3; ++mode) {
    double q[3] = {0.0, 0.0, 0.0};
    if (mine) w_column(wb, mode, q);
    double* o = Y + 9 * rr + 3 * mode;
    o[0] = y[mode][0] + ((d[0] * q[0] + d[1] * q[1]) + d[2] * q[2]);
    o[1] = y[mode][1] + ((d[1] * q[0] + d[3] * q[1]) + d[4] * q[2]);
    o[2] = y[mode][2] + ((d[2] * q[0] + d[4] * q[1]) + d[5] * q[2]);
  }
}

// the three restrictions that follow k_coarse_product: out[a * ostride + ooff + 3 mode + k] = sum over the rows of owned
// aggregate a of (W_i^T Y_i,mode)_k, each by the very sums of k_restrict (same thread / row order, same tree)
__global__ void __launch_bounds__(kThreads, 4) k_restrict_modes(const CoarseView a, const double* __restrict__ Y, double* __restrict__ out,
                                                              int ostride, int ooff) {
  __shared__ double red[8];
  const int g = blockIdx.x;                   // local aggregate
  for (int mode = 0; mode < 3; ++mode) {
    double s0, s1, s2;
    restrict_sums<9>(a, Y + 3 * mode, g, s0, s1, s2);
    const double t0 = block_sum(s0, red);
    const double t1 = block_sum(s1, red);
    const double t2 = block_sum(s2, red);
    if (threadIdx.x == 0) {
      double* o = out + (size_t)(a.a0 + g) * ostride + ooff + 3 * mode;
      o[0] = t0;
      o[1] = t1;
      o[2] = t2;
    }
  }
}

// Dense E (column-major, m = 3 n_agg) from the restricted products, stored aggregate-major so that a rank's share is
// contiguous: R[((a * n_colors + c) * 3 + mode) * 3 + k] is entry (3a+k, 3b+mode) of E0, b = the aggregate of colour c
// next to a; Gm[(a * 3 + mode) * 3 + k] likewise for G = W^T W (block diagonal).  One thread per (a, colour, mode).
__global__ void __launch_bounds__(kThreads) k_coarse_E(const double* __restrict__ R, const double* __restrict__ Gm,
                                                        const int* __restrict__ nbr_of_color, int n_agg, int n_colors,
                                                        double lambda, double* __restrict__ Em) {
  const int t = blockIdx.x * kThreads + threadIdx.x;
  const int total = n_agg * n_colors * 3;
  if (t >= total) return;
  const int mode = t % 3, c = (t / 3) % n_colors, a = t / (3 * n_colors);
  const int b = nbr_of_color[(size_t)a * n_colors + c];
  if (b < 0) return;
  const size_t m = 3 * (size_t)n_agg;
  const size_t col = 3 * (size_t)b + mode;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    double e = R[(((size_t)a * n_colors + c) * 3 + mode) * 3 + k];
    if (a == b) e += lambda * Gm[((size_t)a * 3 + mode) * 3 + k];
    Em[col * m + 3 * (size_t)a + k] = e;
  }
}

__global__ void __launch_bounds__(kThreads) k_set_diagonal(double* __restrict__ Em, int m, double v) {
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i < m) Em[(size_t)i * m + i] = v;
}

// syrk leaves the lower triangle of E^-1: mirror it (column-major, lower -> upper)
__global__ void __launch_bounds__(kThreads) k_symmetrize(double* __restrict__ Em, int m) {
  const size_t t = (size_t)blockIdx.x * kThreads + threadIdx.x;
  if (t >= (size_t)m * m) return;
  const int i = (int)(t % m), j = (int)(t / m);
  if (i < j) Em[t] = Em[(size_t)i * m + j];
}

// mu = Einv c (dense, symmetric, column-major = row-major): kCoarseRows rows per CTA, kThreads / kCoarseRows threads
// per row, every thread eight columns at a time (eight loads in flight), all sums in one fixed order.
constexpr int kCoarseRows = 2;

// Rows [row0, row0 + nrows) of mu = Einv c (this rank's aggregates; one GPU: all of them), the per-row products
// mu_i c_i, and — in the last CTA to finish — their sums over the fixed reduction segments s0 .. s1-1 (segment s =
// rows seg_row[s] .. seg_row[s+1]-1 of the coarse vector: aggregates never cross segment boundaries), left in
// csegs[s0 .. s1).  The segment structure makes c.mu, like every other dot product, independent of the number of
// ranks.  Returns true in every thread of that last CTA.
__device__ __forceinline__ bool coarse_mv_rows(const double* __restrict__ Einv, const double* __restrict__ cvec, int m, int row0,
                                               int nrows, double* __restrict__ mu, double* __restrict__ rowprod,
                                               const int* __restrict__ seg_row, int s0, int s1, double* csegs,
                                               unsigned int* counter, double* red, int* flag) {
  constexpr int kPerRow = kThreads / kCoarseRows;        // threads per row
  const int w = threadIdx.x >> 5;
  const int sub = threadIdx.x / kPerRow, t = threadIdx.x % kPerRow;
  const int lrow = blockIdx.x * kCoarseRows + sub;
  double acc = 0.0;
  if (lrow < nrows) {
    const double* e = Einv + (size_t)(row0 + lrow) * m;
    double p8[8] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    for (int j = t; j < m; j += 8 * kPerRow) {      // the tail is predicated, not a second loop: p8 stays in registers
      double ev[8], cv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int col = j + kPerRow * u;
        const bool in = col < m;
        ev[u] = in ? ldg(e + col) : 0.0;
        cv[u] = in ? ldg(cvec + col) : 0.0;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) p8[u] += ev[u] * cv[u];
    }
    acc = ((p8[0] + p8[1]) + (p8[2] + p8[3])) + ((p8[4] + p8[5]) + (p8[6] + p8[7]));
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[w] = acc;
  __syncthreads();
  if (threadIdx.x < kCoarseRows) {          // thread s finishes row s of this CTA: its warps' sums in warp order
    constexpr int kWarpsPerRow = kPerRow / 32;
    double s = 0.0;
    for (int k = 0; k < kWarpsPerRow; ++k) s += red[threadIdx.x * kWarpsPerRow + k];
    const int rr = blockIdx.x * kCoarseRows + (int)threadIdx.x;
    if (rr < nrows) {
      mu[row0 + rr] = s;
      rowprod[row0 + rr] = s * cvec[row0 + rr];
    }
  }
  if (!last_block(counter, flag)) return false;
  // warp w folds segment s0 + w (, + 8, ...): lane l adds rows l, l + 32, ... of the segment in order, then the butterfly
  const int l = threadIdx.x & 31;
  for (int s = s0 + w; s < s1; s += (int)(blockDim.x >> 5)) {
    double a = 0.0;
    for (int i = seg_row[s] + l; i < seg_row[s + 1]; i += 32) a += __ldcg(rowprod + i);
    a = warp_sum(a);
    if (l == 0) csegs[s] = a;
  }
  __syncthreads();
  return true;
}

// One GPU (and the start of a solve on any number of ranks, where c has been gathered and every rank multiplies all
// rows): mu = Einv c, then the scalar step that closes the preconditioner application,
//     r.z = r.M^-1 r (from the update kernel) + c.mu      — the corrected z is never formed for the dot product.
//   init != 0: start of a solve (r.z of the initial residual, no beta)
__global__ void __launch_bounds__(kThreads, 4) k_coarse_mv(const double* __restrict__ Einv, const double* __restrict__ cvec, int m,
                                                         double* __restrict__ mu, double* __restrict__ rowprod,
                                                         const int* __restrict__ seg_row, unsigned int* counter, Scalars* sc,
                                                         int init) {
  __shared__ double red[8];
  __shared__ double csegs[kSegments];
  __shared__ int flag;
  if (sc->done) return;
  if (!coarse_mv_rows(Einv, cvec, m, 0, m, mu, rowprod, seg_row, 0, kSegments, csegs, counter, red, &flag)) return;
  if (threadIdx.x == 0) {
    double cmu = 0.0;
    for (int s = 0; s < kSegments; ++s) cmu += csegs[s];
    const double rz_new = sc->rz_bj + cmu;
    if (init) {
      sc->rz = rz_new;
      sc->beta = 0.0;
    } else {
      sc->beta = rz_new / sc->rz;
      sc->rz = rz_new;
    }
    sc->rzv[0] = rz_new;   // the peer-memory kernels read r.z of the previous iterate from here
    sc->rzv[1] = rz_new;
    *counter = 0u;
  }
}

// one row of p = (z + W mu) + beta p — ONE spelling for the single-GPU and the peer-memory kernel (same contraction,
// same bits)
__device__ __forceinline__ void dir_coarse_row(const CoarseView& a, const double* __restrict__ z, const double* __restrict__ mu,
                                               double* __restrict__ p, double beta, int r) {
  const int g = a.agg_of[r];
  const WBlock w = w_block(ldg(a.pose + r), ldg(a.ref + g));
  const double m0 = mu[3 * (size_t)g], m1 = mu[3 * (size_t)g + 1], m2 = mu[3 * (size_t)g + 2];
  const size_t o = 3 * (size_t)r;
  const double z0 = z[o] + ((w.c * m0 + w.s * m1) + w.u * m2);
  const double z1 = z[o + 1] + ((w.c * m1 - w.s * m0) + w.v * m2);
  const double z2 = z[o + 2] + m2;
  p[o] = dir_value(z0, beta, p[o]);
  p[o + 1] = dir_value(z1, beta, p[o + 1]);
  p[o + 2] = dir_value(z2, beta, p[o + 2]);
}

// p = (z + W mu) + beta p, one thread per row
__global__ void __launch_bounds__(kThreads) k_dir_coarse(const CoarseView a, const double* __restrict__ z,
                                                          const double* __restrict__ mu, double* __restrict__ p,
                                                          const Scalars* sc) {
  if (sc->done) return;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  if (r >= a.n_own) return;
  dir_coarse_row(a, z, mu, p, sc->beta, r);
}

}  // namespace gs
