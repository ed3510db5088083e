/* CPU PORT of the pose-graph optimisation path (C99 + OpenMP).  TEST INFRASTRUCTURE / TIMED BASELINE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this;
 * the product (libgraphslam_cuda) never links or calls it.
 *
 * PARITY UNPINNED: like oracle/gs_oracle.py this restates GTSAM 4.0.x semantics (the reference's arithmetic
 * lives in an un-vendored, un-pinned GTSAM: reference src/graph/CMakeLists.txt:9,18) and is pinned by our
 * own known-answer tests plus agreement with gs_oracle.py (exact sparse-LU solve) to 1e-9.
 *
 *   reference call site                                   restated here
 *   graph.cpp:87-92,106-111 BetweenFactor<Pose2>          lin_between()
 *   graph.cpp:57            PriorFactor<Pose2>            lin_prior()
 *   graph.cpp:45,83,103     Gaussian::Covariance(Q)       information()
 *   graph.cpp:119           LevenbergMarquardtOptimizer(graph, initial).optimize()   gs_cpu_optimize()
 *   graph.cpp:122-125       x(), y(), theta()             theta = atan2(s, c)
 *
 * Linear solve, two variants (cpu_params.solver):
 *   0  block-Jacobi PCG on the 3x3-block CSR normal equations (the same algorithm the CUDA path runs, so iteration
 *      counts are comparable), all host threads via OpenMP;
 *   1  sparse multifrontal Cholesky with a nested-dissection ordering (oracle/cpu_chol.c): what GTSAM's default
 *      MULTIFRONTAL_CHOLESKY does behind graph.cpp:119, and an exact solve written independently of gs_oracle.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "cpu_chol.h"

typedef struct {
  int max_iterations;
  double rel_error_tol, abs_error_tol, error_tol;
  double lambda_initial, lambda_factor, lambda_upper, lambda_lower, min_model_fidelity;
  int diagonal_damping, jacobian_mode;
  double pcg_rel_tol;
  int pcg_max_iters;
  int pcg_iter_cap_total; /* >0: stop after this many PCG iterations in total (bounded timing sample) */
  double pcg_time_cap_ms; /* >0: a PCG solve gives up after this much wall time (bounded timing sample) */
  int solver;             /* 0: block-Jacobi PCG (the CUDA path's algorithm); 1: sparse multifrontal Cholesky (what the
                             reference's GTSAM call does, graph.cpp:119; oracle/cpu_chol.c) */
} cpu_params;

typedef struct {
  int iterations, outer_loops, inner_iterations, converged;
  double chi2_initial, chi2_final, lambda_final;
  int64_t pcg_iterations_total;
  double ms_structure, ms_linearize, ms_solve, ms_update, ms_total;
  int threads;
  double ms_analyze, ms_factor;     /* solver 1: ordering + symbolic (once), numeric factorisations (summed) */
  double chol_nnz, chol_flops;      /* solver 1: entries of L, flops of one factorisation */
  int capped;                       /* 1: a PCG cap (iterations or time) ended the run before the tolerance */
} cpu_stats;

static double now_ms(void) {
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

void gs_cpu_default_params(cpu_params* p) {
  memset(p, 0, sizeof *p);
  p->max_iterations = 100;
  p->rel_error_tol = 1e-5;
  p->abs_error_tol = 1e-5;
  p->error_tol = 0.0;
  p->lambda_initial = 1e-5;
  p->lambda_factor = 10.0;
  p->lambda_upper = 1e5;
  p->lambda_lower = 0.0;
  p->min_model_fidelity = 1e-3;
  p->pcg_rel_tol = 1e-12;
  p->pcg_max_iters = 100000;
}

int gs_cpu_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* explicit thread count (torchrun exports OMP_NUM_THREADS=1 to its children; a timed baseline must not inherit that) */
void gs_cpu_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* Omega = Q^-1 (symmetric 6: 00 01 02 11 12 22); returns 0 if Q is not SPD */
static int information(const double* q, double* o) {
  const double a00 = q[0], a11 = q[4], a22 = q[8];
  const double a01 = 0.5 * (q[1] + q[3]), a02 = 0.5 * (q[2] + q[6]), a12 = 0.5 * (q[5] + q[7]);
  if (!(a00 > 0)) return 0;
  const double l10 = a01 / sqrt(a00), l20 = a02 / sqrt(a00);
  const double d1 = a11 - l10 * l10;
  if (!(d1 > 0)) return 0;
  const double l21 = (a12 - l20 * l10) / sqrt(d1);
  if (!(a22 - l20 * l20 - l21 * l21 > 0)) return 0;
  const double c00 = a11 * a22 - a12 * a12, c01 = a02 * a12 - a01 * a22, c02 = a01 * a12 - a02 * a11;
  const double id = 1.0 / (a00 * c00 + a01 * c01 + a02 * c02);
  o[0] = c00 * id; o[1] = c01 * id; o[2] = c02 * id;
  o[3] = (a00 * a22 - a02 * a02) * id; o[4] = (a01 * a02 - a00 * a12) * id; o[5] = (a00 * a11 - a01 * a01) * id;
  return 1;
}

/* BetweenFactor<Pose2>: e = Local(z^-1 o (xi^-1 o xj)); Ai = H1 (or D*H1), Aj = I (or D). poses are (x,y,c,s) */
static void lin_between(const double* pi, const double* pj, const double* z, const double* zcs, int exact, double* e,
                        double* Ai, double* Aj) {
  const double ci = pi[2], si = pi[3], cj = pj[2], sj = pj[3];
  const double dx = pj[0] - pi[0], dy = pj[1] - pi[1];
  const double hx = ci * dx + si * dy, hy = -si * dx + ci * dy;
  double c = ci * cj + si * sj, s = -si * cj + ci * sj;
  { /* gtsam::Rot2::normalize(): only when c^2 + s^2 is off 1 by more than 1e-10 */
    const double scale = c * c + s * s;
    if (fabs(scale - 1.0) > 1e-10) { const double inv = 1.0 / sqrt(scale); c *= inv; s *= inv; }
  }
  const double cm = zcs[0], sm = zcs[1];
  const double rx = hx - z[0], ry = hy - z[1];
  e[0] = cm * rx + sm * ry;
  e[1] = -sm * rx + cm * ry;
  const double se = s * cm - c * sm, ce = c * cm + s * sm;
  e[2] = atan2(se, ce);
  if (!Ai) return;
  const double dt1 = -sj * dx + cj * dy, dt2 = -cj * dx - sj * dy;
  double H[9] = {-c, -s, dt1, s, -c, dt2, 0, 0, -1};
  if (!exact) {
    memcpy(Ai, H, sizeof H);
    const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    memcpy(Aj, I, sizeof I);
  } else {
    const double n = 1.0 / sqrt(ce * ce + se * se), cE = ce * n, sE = se * n;
    for (int k = 0; k < 3; ++k) {
      Ai[k] = cE * H[k] - sE * H[3 + k];
      Ai[3 + k] = sE * H[k] + cE * H[3 + k];
      Ai[6 + k] = H[6 + k];
    }
    const double D[9] = {cE, -sE, 0, sE, cE, 0, 0, 0, 1};
    memcpy(Aj, D, sizeof D);
  }
}

/* PriorFactor<Pose2> (GTSAM 4.0): e = -Local(x^-1 o prior), H = I */
static void lin_prior(const double* p, const double* pr, int exact, double* e, double* A) {
  const double c = p[2], s = p[3];
  const double dx = pr[0] - p[0], dy = pr[1] - p[1];
  e[0] = -(c * dx + s * dy);
  e[1] = -(-s * dx + c * dy);
  e[2] = -atan2(pr[3] * c - pr[2] * s, pr[2] * c + pr[3] * s);
  if (!A) return;
  const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  memcpy(A, I, sizeof I);
  if (exact) { A[2] = e[1]; A[5] = -e[0]; }
}

static double quad(const double* o, const double* e) {
  const double w0 = o[0] * e[0] + o[1] * e[1] + o[2] * e[2];
  const double w1 = o[1] * e[0] + o[3] * e[1] + o[4] * e[2];
  const double w2 = o[2] * e[0] + o[4] * e[1] + o[5] * e[2];
  return e[0] * w0 + e[1] * w1 + e[2] * w2;
}

/* out(3x3) += Ar^T Omega Ac ; g(3) += Ar^T Omega e */
static void add_pieces(const double* o, const double* Ar, const double* Ac, const double* e, double* out, double* g) {
  double W[9];
  for (int c = 0; c < 3; ++c) {
    W[c] = o[0] * Ar[c] + o[1] * Ar[3 + c] + o[2] * Ar[6 + c];
    W[3 + c] = o[1] * Ar[c] + o[3] * Ar[3 + c] + o[4] * Ar[6 + c];
    W[6 + c] = o[2] * Ar[c] + o[4] * Ar[3 + c] + o[5] * Ar[6 + c];
  }
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) out[3 * r + c] += W[r] * Ac[c] + W[3 + r] * Ac[3 + c] + W[6 + r] * Ac[6 + c];
    if (g) g[r] += W[r] * e[0] + W[3 + r] * e[1] + W[6 + r] * e[2];
  }
}

typedef struct {
  int n, e, np;
  const int *ei, *ej, *pidx;
  const double* z;
  double *zcs, *om, *ppose, *pom;
  /* block CSR incl. diagonal */
  int *rowptr, *colidx, *diag;
  /* incidence: per row the (edge<<1|side) list, sorted by column */
  int *irowptr, *inc, *incblk;
  int *prowptr, *pinc;
  double *H, *b;
} problem;

static int cmp_pair(const void* a, const void* b) {
  const int64_t x = *(const int64_t*)a, y = *(const int64_t*)b;
  return (x > y) - (x < y);
}

static int build(problem* P) {
  const int n = P->n, e = P->e;
  P->irowptr = (int*)calloc(n + 1, sizeof(int));
  P->prowptr = (int*)calloc(n + 1, sizeof(int));
  for (int k = 0; k < e; ++k) { P->irowptr[P->ei[k] + 1]++; P->irowptr[P->ej[k] + 1]++; }
  for (int k = 0; k < P->np; ++k) P->prowptr[P->pidx[k] + 1]++;
  for (int r = 0; r < n; ++r) { P->irowptr[r + 1] += P->irowptr[r]; P->prowptr[r + 1] += P->prowptr[r]; }
  int64_t* tmp = (int64_t*)malloc(sizeof(int64_t) * 2 * (size_t)(e > 0 ? e : 1));
  int* fill = (int*)malloc(sizeof(int) * (n + 1));
  memcpy(fill, P->irowptr, sizeof(int) * (n + 1));
  /* key = (col << 32) | (edge<<1|side): sorting a row's keys orders by column then edge */
  for (int k = 0; k < e; ++k) {
    tmp[fill[P->ei[k]]++] = ((int64_t)P->ej[k] << 32) | (uint32_t)(k << 1);
    tmp[fill[P->ej[k]]++] = ((int64_t)P->ei[k] << 32) | (uint32_t)((k << 1) | 1);
  }
  P->inc = (int*)malloc(sizeof(int) * 2 * (size_t)(e > 0 ? e : 1));
  P->incblk = (int*)malloc(sizeof(int) * 2 * (size_t)(e > 0 ? e : 1));
  P->rowptr = (int*)calloc(n + 1, sizeof(int));
#pragma omp parallel for schedule(dynamic, 1024)
  for (int r = 0; r < n; ++r) {
    const int a = P->irowptr[r], b = P->irowptr[r + 1];
    qsort(tmp + a, b - a, sizeof(int64_t), cmp_pair);
    int cnt = 1, last = -1; /* diagonal */
    for (int k = a; k < b; ++k) {
      const int col = (int)(tmp[k] >> 32);
      if (col != last) { ++cnt; last = col; }
    }
    P->rowptr[r + 1] = cnt;
  }
  for (int r = 0; r < n; ++r) P->rowptr[r + 1] += P->rowptr[r];
  const int nnzb = P->rowptr[n];
  P->colidx = (int*)malloc(sizeof(int) * (size_t)nnzb);
  P->diag = (int*)malloc(sizeof(int) * (size_t)n);
#pragma omp parallel for schedule(dynamic, 1024)
  for (int r = 0; r < n; ++r) {
    const int a = P->irowptr[r], b = P->irowptr[r + 1];
    int w = P->rowptr[r], last = -1, placed = 0;
    for (int k = a; k < b; ++k) {
      const int col = (int)(tmp[k] >> 32);
      if (!placed && col > r) { P->diag[r] = w; P->colidx[w++] = r; placed = 1; }
      if (col != last) { P->colidx[w++] = col; last = col; }
      P->inc[k] = (int)(uint32_t)tmp[k];
      P->incblk[k] = w - 1;
    }
    if (!placed) { P->diag[r] = w; P->colidx[w++] = r; }
  }
  memcpy(fill, P->prowptr, sizeof(int) * (n + 1));
  P->pinc = (int*)malloc(sizeof(int) * (size_t)(P->np > 0 ? P->np : 1));
  for (int k = 0; k < P->np; ++k) P->pinc[fill[P->pidx[k]]++] = k;
  free(tmp);
  free(fill);
  P->H = (double*)malloc(sizeof(double) * 9 * (size_t)nnzb);
  P->b = (double*)malloc(sizeof(double) * 3 * (size_t)n);
  return nnzb;
}

static void destroy(problem* P) {
  free(P->irowptr); free(P->prowptr); free(P->inc); free(P->incblk); free(P->pinc);
  free(P->rowptr); free(P->colidx); free(P->diag); free(P->H); free(P->b);
  free(P->zcs); free(P->om); free(P->ppose); free(P->pom);
}

/* graph.error(values)*2 */
static double chi2_eval(const problem* P, const double* pose) {
  double tot = 0.0;
#pragma omp parallel for reduction(+ : tot) schedule(static)
  for (int k = 0; k < P->e; ++k) {
    double e[3];
    lin_between(pose + 4 * (size_t)P->ei[k], pose + 4 * (size_t)P->ej[k], P->z + 3 * (size_t)k, P->zcs + 2 * (size_t)k, 0, e, 0, 0);
    tot += quad(P->om + 6 * (size_t)k, e);
  }
  for (int k = 0; k < P->np; ++k) {
    double e[3];
    lin_prior(pose + 4 * (size_t)P->pidx[k], P->ppose + 4 * (size_t)k, 0, e, 0);
    tot += quad(P->pom + 6 * (size_t)k, e);
  }
  return tot;
}

/* row-owner assembly: every output block is written by the thread that owns its row */
static void linearize(problem* P, const double* pose, int exact) {
  const int n = P->n;
#pragma omp parallel for schedule(dynamic, 512)
  for (int r = 0; r < n; ++r) {
    double* Hr = P->H + 9 * (size_t)P->rowptr[r];
    memset(Hr, 0, sizeof(double) * 9 * (size_t)(P->rowptr[r + 1] - P->rowptr[r]));
    double* D = P->H + 9 * (size_t)P->diag[r];
    double g[3] = {0, 0, 0};
    for (int q = P->prowptr[r]; q < P->prowptr[r + 1]; ++q) {
      const int k = P->pinc[q];
      double e[3], A[9];
      lin_prior(pose + 4 * (size_t)r, P->ppose + 4 * (size_t)k, exact, e, A);
      add_pieces(P->pom + 6 * (size_t)k, A, A, e, D, g);
    }
    for (int q = P->irowptr[r]; q < P->irowptr[r + 1]; ++q) {
      const int k = P->inc[q] >> 1, side = P->inc[q] & 1;
      double e[3], Ai[9], Aj[9];
      lin_between(pose + 4 * (size_t)P->ei[k], pose + 4 * (size_t)P->ej[k], P->z + 3 * (size_t)k, P->zcs + 2 * (size_t)k, exact, e, Ai, Aj);
      const double* Ar = side ? Aj : Ai;
      const double* Ac = side ? Ai : Aj;
      const double* o = P->om + 6 * (size_t)k;
      add_pieces(o, Ar, Ar, e, D, g);
      add_pieces(o, Ar, Ac, e, P->H + 9 * (size_t)P->incblk[q], 0);
    }
    P->b[3 * (size_t)r] = g[0]; P->b[3 * (size_t)r + 1] = g[1]; P->b[3 * (size_t)r + 2] = g[2];
  }
}

/* y = (H + lambda*diag(d)) p ; returns p.y */
static double spmv(const problem* P, const double* p, double* y, double lambda, const double* dscale) {
  double dot = 0.0;
#pragma omp parallel for reduction(+ : dot) schedule(static)
  for (int r = 0; r < P->n; ++r) {
    double a0 = 0, a1 = 0, a2 = 0;
    for (int k = P->rowptr[r]; k < P->rowptr[r + 1]; ++k) {
      const double* h = P->H + 9 * (size_t)k;
      const double* x = p + 3 * (size_t)P->colidx[k];
      a0 += h[0] * x[0] + h[1] * x[1] + h[2] * x[2];
      a1 += h[3] * x[0] + h[4] * x[1] + h[5] * x[2];
      a2 += h[6] * x[0] + h[7] * x[1] + h[8] * x[2];
    }
    const double* x = p + 3 * (size_t)r;
    if (dscale) {
      a0 += lambda * dscale[3 * (size_t)r] * x[0]; a1 += lambda * dscale[3 * (size_t)r + 1] * x[1]; a2 += lambda * dscale[3 * (size_t)r + 2] * x[2];
    } else {
      a0 += lambda * x[0]; a1 += lambda * x[1]; a2 += lambda * x[2];
    }
    y[3 * (size_t)r] = a0; y[3 * (size_t)r + 1] = a1; y[3 * (size_t)r + 2] = a2;
    dot += x[0] * a0 + x[1] * a1 + x[2] * a2;
  }
  return dot;
}

static void build_minv(const problem* P, double lambda, const double* dscale, double* M) {
#pragma omp parallel for schedule(static)
  for (int r = 0; r < P->n; ++r) {
    const double* d = P->H + 9 * (size_t)P->diag[r];
    const double l0 = dscale ? lambda * dscale[3 * (size_t)r] : lambda;
    const double l1 = dscale ? lambda * dscale[3 * (size_t)r + 1] : lambda;
    const double l2 = dscale ? lambda * dscale[3 * (size_t)r + 2] : lambda;
    const double a00 = d[0] + l0, a01 = d[1], a02 = d[2], a11 = d[4] + l1, a12 = d[5], a22 = d[8] + l2;
    const double c00 = a11 * a22 - a12 * a12, c01 = a02 * a12 - a01 * a22, c02 = a01 * a12 - a02 * a11;
    const double id = 1.0 / (a00 * c00 + a01 * c01 + a02 * c02);
    double* m = M + 6 * (size_t)r;
    m[0] = c00 * id; m[1] = c01 * id; m[2] = c02 * id;
    m[3] = (a00 * a22 - a02 * a02) * id; m[4] = (a01 * a02 - a00 * a12) * id; m[5] = (a00 * a11 - a01 * a01) * id;
  }
}

/* block-Jacobi PCG for (H + lambda D) x = -b ; returns iterations, -1 on breakdown */
static int pcg(const problem* P, double lambda, const double* dscale, const double* M, double* x, double* r, double* z,
               double* p, double* Ap, double tol, int max_iters, double time_cap_ms, double* relres) {
  const int n = P->n;
  const double t_start = now_ms();
  double rz = 0, rr0 = 0;
#pragma omp parallel for reduction(+ : rz, rr0) schedule(static)
  for (int v = 0; v < n; ++v) {
    const size_t o = 3 * (size_t)v;
    const double* m = M + 6 * (size_t)v;
    const double r0 = -P->b[o], r1 = -P->b[o + 1], r2 = -P->b[o + 2];
    const double z0 = m[0] * r0 + m[1] * r1 + m[2] * r2, z1 = m[1] * r0 + m[3] * r1 + m[4] * r2, z2 = m[2] * r0 + m[4] * r1 + m[5] * r2;
    x[o] = x[o + 1] = x[o + 2] = 0.0;
    r[o] = r0; r[o + 1] = r1; r[o + 2] = r2;
    z[o] = p[o] = z0; z[o + 1] = p[o + 1] = z1; z[o + 2] = p[o + 2] = z2;
    rz += r0 * z0 + r1 * z1 + r2 * z2;
    rr0 += r0 * r0 + r1 * r1 + r2 * r2;
  }
  *relres = 0.0;
  if (rr0 == 0.0) return 0;
  int it = 0;
  double rr = rr0;
  while (it < max_iters) {
    const double pAp = spmv(P, p, Ap, lambda, dscale);
    if (!(pAp > 0.0)) { *relres = sqrt(rr / rr0); return -1; }
    const double alpha = rz / pAp;
    double rzn = 0;
    rr = 0;
#pragma omp parallel for reduction(+ : rzn, rr) schedule(static)
    for (int v = 0; v < n; ++v) {
      const size_t o = 3 * (size_t)v;
      const double* m = M + 6 * (size_t)v;
      x[o] += alpha * p[o]; x[o + 1] += alpha * p[o + 1]; x[o + 2] += alpha * p[o + 2];
      const double r0 = r[o] - alpha * Ap[o], r1 = r[o + 1] - alpha * Ap[o + 1], r2 = r[o + 2] - alpha * Ap[o + 2];
      const double z0 = m[0] * r0 + m[1] * r1 + m[2] * r2, z1 = m[1] * r0 + m[3] * r1 + m[4] * r2, z2 = m[2] * r0 + m[4] * r1 + m[5] * r2;
      r[o] = r0; r[o + 1] = r1; r[o + 2] = r2;
      z[o] = z0; z[o + 1] = z1; z[o + 2] = z2;
      rzn += r0 * z0 + r1 * z1 + r2 * z2;
      rr += r0 * r0 + r1 * r1 + r2 * r2;
    }
    ++it;
    if (rr <= tol * tol * rr0) break;
    if (time_cap_ms > 0.0 && (it & 15) == 0 && now_ms() - t_start > time_cap_ms) break;
    const double beta = rzn / rz;
    rz = rzn;
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < 3 * (size_t)n; ++i) p[i] = z[i] + beta * p[i];
  }
  *relres = sqrt(rr / rr0);
  return it;
}

/* LevenbergMarquardtOptimizer(graph, initial).optimize() — poses_xyt (n*3) in/out, dense indices.
 * trace_err[cap+1], trace_tries[cap] optional.  Returns 0 ok, 2 not SPD, 7 PCG not converged. */
int gs_cpu_optimize(int n, int e, const int* ei, const int* ej, const double* z, const double* cov9, int np, const int* pidx,
                    const double* ppose_xyt, const double* pcov9, double* poses_xyt, const cpu_params* prm, cpu_stats* st,
                    double* trace_err, int* trace_tries, int trace_cap) {
  const double t_begin = now_ms();
  cpu_params P_;
  if (prm) P_ = *prm; else gs_cpu_default_params(&P_);
  const cpu_params* p = &P_;
  problem G;
  memset(&G, 0, sizeof G);
  G.n = n; G.e = e; G.np = np; G.ei = ei; G.ej = ej; G.pidx = pidx; G.z = z;
  G.zcs = (double*)malloc(sizeof(double) * 2 * (size_t)(e > 0 ? e : 1));
  G.om = (double*)malloc(sizeof(double) * 6 * (size_t)(e > 0 ? e : 1));
  G.ppose = (double*)malloc(sizeof(double) * 4 * (size_t)(np > 0 ? np : 1));
  G.pom = (double*)malloc(sizeof(double) * 6 * (size_t)(np > 0 ? np : 1));
  int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
  for (int k = 0; k < e; ++k) {
    G.zcs[2 * (size_t)k] = cos(z[3 * (size_t)k + 2]);
    G.zcs[2 * (size_t)k + 1] = sin(z[3 * (size_t)k + 2]);
    if (!information(cov9 + 9 * (size_t)k, G.om + 6 * (size_t)k)) bad |= 1;
  }
  for (int k = 0; k < np; ++k) {
    G.ppose[4 * k] = ppose_xyt[3 * k]; G.ppose[4 * k + 1] = ppose_xyt[3 * k + 1];
    G.ppose[4 * k + 2] = cos(ppose_xyt[3 * k + 2]); G.ppose[4 * k + 3] = sin(ppose_xyt[3 * k + 2]);
    if (!information(pcov9 + 9 * (size_t)k, G.pom + 6 * (size_t)k)) bad |= 1;
  }
  if (bad) { free(G.zcs); free(G.om); free(G.ppose); free(G.pom); return 2; }
  build(&G);
  cpu_stats S;
  memset(&S, 0, sizeof S);
  S.threads = gs_cpu_threads();
  S.ms_structure = now_ms() - t_begin;
  chol* CH = 0;
  if (p->solver == 1) {
    const double ta = now_ms();
    CH = chol_analyze(n, G.rowptr, G.colidx);
    S.ms_analyze = now_ms() - ta;
    chol_info(CH, &S.chol_nnz, &S.chol_flops, 0, 0);
  }
  const size_t n3 = 3 * (size_t)n;
  double* pose = (double*)malloc(sizeof(double) * 4 * (size_t)n);
  double* cand = (double*)malloc(sizeof(double) * 4 * (size_t)n);
  double* buf = (double*)malloc(sizeof(double) * (5 * n3 + 6 * (size_t)n + n3));
  double *x = buf, *r = x + n3, *zv = r + n3, *pv = zv + n3, *Ap = pv + n3, *M = Ap + n3, *dscale = M + 6 * (size_t)n;
#pragma omp parallel for schedule(static)
  for (int v = 0; v < n; ++v) {
    pose[4 * (size_t)v] = poses_xyt[3 * (size_t)v]; pose[4 * (size_t)v + 1] = poses_xyt[3 * (size_t)v + 1];
    pose[4 * (size_t)v + 2] = cos(poses_xyt[3 * (size_t)v + 2]); pose[4 * (size_t)v + 3] = sin(poses_xyt[3 * (size_t)v + 2]);
  }
  double error = 0.5 * chi2_eval(&G, pose);
  S.chi2_initial = 2 * error;
  if (trace_err && trace_cap >= 0) trace_err[0] = error;
  double lambda = p->lambda_initial;
  int iterations = 0, rc = 0, ntrace = 0, capped = 0;
  /* NonlinearOptimizer::defaultOptimize: nothing to do when the initial error is already <= errorTol */
  if (error <= p->error_tol) S.converged = 1;
  while (error > p->error_tol && iterations < p->max_iterations && !capped) {
    const double current_error = error;
    double t0 = now_ms();
    linearize(&G, pose, p->jacobian_mode);
    if (p->diagonal_damping) {
#pragma omp parallel for schedule(static)
      for (int v = 0; v < n; ++v) {
        const double* d = G.H + 9 * (size_t)G.diag[v];
        dscale[3 * (size_t)v] = fmin(fmax(d[0], 1e-6), 1e32);
        dscale[3 * (size_t)v + 1] = fmin(fmax(d[4], 1e-6), 1e32);
        dscale[3 * (size_t)v + 2] = fmin(fmax(d[8], 1e-6), 1e32);
      }
    }
    const double* ds = p->diagonal_damping ? dscale : 0;
    S.ms_linearize += now_ms() - t0;
    S.outer_loops++;
    int tries = 0;
    for (;;) {
      ++tries;
      S.inner_iterations++;
      t0 = now_ms();
      double relres = 0.0;
      int it;
      if (CH) {   /* exact solve: (H + lambda D) x = -b */
        const int bad = chol_factor(CH, G.H, G.diag, lambda, ds);
        S.ms_factor += now_ms() - t0;
        if (!bad) {
#pragma omp parallel for schedule(static)
          for (size_t i = 0; i < n3; ++i) r[i] = -G.b[i];
          chol_solve(CH, r, x);
        }
        it = bad ? -1 : 0;
      } else {
        build_minv(&G, lambda, ds, M);
        int max_it = p->pcg_max_iters;
        if (p->pcg_iter_cap_total > 0) {
          const int64_t left = (int64_t)p->pcg_iter_cap_total - S.pcg_iterations_total;
          if (left < max_it) max_it = (int)(left > 0 ? left : 0);
        }
        it = pcg(&G, lambda, ds, M, x, r, zv, pv, Ap, p->pcg_rel_tol, max_it, p->pcg_time_cap_ms, &relres);
        if (p->pcg_time_cap_ms > 0.0 && relres > p->pcg_rel_tol) capped = 1;
      }
      S.ms_solve += now_ms() - t0;
      const int solved = it >= 0;
      if (it < 0) it = 0;
      S.pcg_iterations_total += it;
      if (p->pcg_iter_cap_total > 0 && S.pcg_iterations_total >= p->pcg_iter_cap_total) capped = 1;
      if (solved && relres > p->pcg_rel_tol && !capped) rc = 7;
      int step_ok = 0, stop_search = 0;
      double new_error = INFINITY;
      t0 = now_ms();
      if (solved) {
        const double dHd = spmv(&G, x, Ap, 0.0, 0);
        double db = 0;
#pragma omp parallel for reduction(+ : db) schedule(static)
        for (size_t i = 0; i < n3; ++i) db += x[i] * G.b[i];
        const double lin_change = -db - 0.5 * dHd;
        if (lin_change >= 0.0) {
#pragma omp parallel for schedule(static)
          for (int v = 0; v < n; ++v) { /* Values::retract */
            const double* q = pose + 4 * (size_t)v;
            const double* d = x + 3 * (size_t)v;
            const double cd = cos(d[2]), sd = sin(d[2]);
            double* o = cand + 4 * (size_t)v;
            o[0] = q[0] + q[2] * d[0] - q[3] * d[1];
            o[1] = q[1] + q[3] * d[0] + q[2] * d[1];
            const double c = q[2] * cd - q[3] * sd, s = q[3] * cd + q[2] * sd, inv = 1.0 / sqrt(c * c + s * s);
            o[2] = c * inv; o[3] = s * inv;
          }
          new_error = 0.5 * chi2_eval(&G, cand);
          const double cost_change = error - new_error;
          if (lin_change > 2.220446049250313e-16 * error) step_ok = (cost_change / lin_change) > p->min_model_fidelity;
          if (fabs(cost_change) < p->rel_error_tol * error) stop_search = 1;
        }
      }
      S.ms_update += now_ms() - t0;
      if (capped) break;
      if (step_ok) {
        double* t = pose; pose = cand; cand = t;
        error = new_error;
        lambda = fmax(p->lambda_lower, lambda / p->lambda_factor);
        ++iterations;
        break;
      } else if (!stop_search) {
        lambda *= p->lambda_factor;
        if (lambda >= p->lambda_upper) break;
      } else {
        break;
      }
    }
    if (ntrace < trace_cap) {
      if (trace_tries) trace_tries[ntrace] = tries;
      if (trace_err) trace_err[ntrace + 1] = error;
    }
    ++ntrace;
    if (iterations >= p->max_iterations) break;
    if (error <= p->error_tol) { S.converged = 1; break; }
    const double abs_dec = current_error - error, rel_dec = current_error != 0 ? abs_dec / current_error : 0;
    if ((p->rel_error_tol != 0 && rel_dec <= p->rel_error_tol) || abs_dec <= p->abs_error_tol) { S.converged = 1; break; }
    if (!isfinite(current_error)) break;
  }
#pragma omp parallel for schedule(static)
  for (int v = 0; v < n; ++v) {
    poses_xyt[3 * (size_t)v] = pose[4 * (size_t)v];
    poses_xyt[3 * (size_t)v + 1] = pose[4 * (size_t)v + 1];
    poses_xyt[3 * (size_t)v + 2] = atan2(pose[4 * (size_t)v + 3], pose[4 * (size_t)v + 2]);
  }
  S.iterations = iterations;
  S.capped = capped;
  S.chi2_final = 2 * error;
  S.lambda_final = lambda;
  S.ms_total = now_ms() - t_begin;
  if (st) *st = S;
  free(pose); free(cand); free(buf);
  chol_free(CH);
  destroy(&G);
  return rc;
}
