/* TEST INFRASTRUCTURE — never shipped, never linked into the product (see oracle/README in cpu_ref.c's header).
 *
 * Sparse direct solver of the damped normal equations (H + lambda D) x = g for the CPU restatement of the path:
 * what the reference actually runs at /root/reference/src/graph/src/graph.cpp:119 is GTSAM's
 * LevenbergMarquardtOptimizer with its default linear solver, MULTIFRONTAL_CHOLESKY (GTSAM
 * NonlinearOptimizerParams::linearSolverType default).  GTSAM is absent from this image and from
 * /root/reference, so this file restates the published algorithm:
 *
 *   ordering      nested dissection of the block graph of H (George 1973 "automatic nested dissection": level
 *                 structure rooted at a pseudo-peripheral vertex, the middle level is the separator) — the role
 *                 COLAMD / METIS play in GTSAM's Ordering;
 *   symbolic      one front per separator / leaf: its pivot vertices and the boundary vertices (in ancestors'
 *                 separators) its subtree touches — GTSAM's Bayes-tree cliques (frontals + separator);
 *   numeric       multifrontal Cholesky (Duff & Reid 1983): assemble the front from H and the children's update
 *                 matrices, dense partial Cholesky of the pivot columns, Schur complement to the parent —
 *                 GTSAM's EliminateCholesky over the clique tree;
 *   solve         forward and backward substitution over the tree.
 *
 * Unknowns are 3x3 blocks (SE(2) poses); fronts are dense column-major scalar matrices.  OpenMP: independent
 * subtrees are tasks, the trailing update of a large front is a taskloop.  Exact up to rounding, so the LM driver
 * of cpu_ref.c run with solver = 1 is an exact-solve restatement written independently of oracle/gs_oracle.py
 * (SuperLU): tests/test_cpu_port.py pins the two against each other.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "cpu_chol.h"

#define LEAF_MAX 48      /* vertices per leaf front */
#define NB 96            /* pivot columns per dense panel (multiple of 3 and of NR) */
#define MR 8
#define NR 6

typedef struct {
  int nv, nb;            /* pivot vertices, boundary vertices */
  int v0;                /* elimination position of the first pivot vertex */
  int parent;
  int nchild, child0;    /* children: ids child_list[child0 .. child0 + nchild) */
  int* verts;            /* pivot vertices (original ids); freed after positions are assigned */
  int* bnd;              /* boundary: elimination positions, ascending */
  int* rel;              /* per boundary entry: row block in the parent's front */
  double* L;             /* (3(nv+nb)) x (3 nv) column-major: [L11; L21] */
  double* U;             /* update matrix 3nb x 3nb (lower), alive between this node's and its parent's factorisation */
  double* y;             /* solve workspace: 3(nv+nb) */
} node_t;

struct chol {
  int n;
  const int *rowptr, *colidx;
  int *perm, *iperm;     /* perm[pos] = vertex, iperm[vertex] = pos */
  node_t* nodes;
  int nnodes, cap;
  int* child_list;
  int* roots;
  int nroots;
  int* post;             /* node ids, children before parents */
  double nnz_l, flops;
  int max_front;
  int failed;
};

/* ------------------------------------------------------------------------------------------- ordering */
typedef struct {
  chol* C;
  int* level;            /* scratch per vertex */
  int* region;           /* region id per vertex; -1 = already placed in a separator / leaf */
  int next_region;
} nd_ctx;

static int new_node(chol* C, int nv, const int* verts) {
  int id;
#pragma omp critical(chol_nodes)
  {
    if (C->nnodes == C->cap) {
      C->cap = C->cap ? 2 * C->cap : 1024;
      C->nodes = (node_t*)realloc(C->nodes, sizeof(node_t) * (size_t)C->cap);
    }
    id = C->nnodes++;
    memset(&C->nodes[id], 0, sizeof(node_t));
    C->nodes[id].parent = -1;
  }
  int* v = (int*)malloc(sizeof(int) * (size_t)(nv > 0 ? nv : 1));
  memcpy(v, verts, sizeof(int) * (size_t)nv);
#pragma omp critical(chol_nodes)
  {
    C->nodes[id].nv = nv;
    C->nodes[id].verts = v;
  }
  return id;
}

static int new_region(nd_ctx* X) {
  int r;
#pragma omp atomic capture
  r = X->next_region++;
  return r;
}

/* BFS inside region `rid` from `root`; queue receives the visit order, level[] the depth; returns count, *nlev */
static int bfs(nd_ctx* X, int rid, int root, int* queue, int* nlev) {
  const int *rp = X->C->rowptr, *ci = X->C->colidx;
  int head = 0, tail = 0, maxl = 0;
  queue[tail++] = root;
  X->level[root] = 0;
  /* visited marker: level >= 0 ; caller resets level to -1 for the region before calling */
  while (head < tail) {
    const int v = queue[head++];
    const int lv = X->level[v];
    if (lv > maxl) maxl = lv;
    for (int k = rp[v]; k < rp[v + 1]; ++k) {
      const int u = ci[k];
      if (u == v || X->region[u] != rid || X->level[u] >= 0) continue;
      X->level[u] = lv + 1;
      queue[tail++] = u;
    }
  }
  *nlev = maxl + 1;
  return tail;
}

typedef struct { int* ids; int n, cap; } ilist;
static void ilist_push(ilist* l, int v) {
  if (l->n == l->cap) {
    l->cap = l->cap ? 2 * l->cap : 8;
    l->ids = (int*)realloc(l->ids, sizeof(int) * (size_t)l->cap);
  }
  l->ids[l->n++] = v;
}

static void nd_region(nd_ctx* X, int* verts, int cnt, int rid, ilist* out);

/* one connected component (all of region rid): returns the node that heads its elimination subtree */
static int nd_component(nd_ctx* X, int* verts, int cnt, int rid) {
  chol* C = X->C;
  if (cnt <= LEAF_MAX) {
    for (int k = 0; k < cnt; ++k) X->region[verts[k]] = -1;
    return new_node(C, cnt, verts);
  }
  int* queue = (int*)malloc(sizeof(int) * (size_t)cnt);
  int nlev = 0;
  /* pseudo-peripheral root: two sweeps, restarting from a vertex of the last level */
  int root = verts[0];
  for (int sweep = 0; sweep < 2; ++sweep) {
    for (int k = 0; k < cnt; ++k) X->level[verts[k]] = -1;
    bfs(X, rid, root, queue, &nlev);
    root = queue[cnt - 1];
  }
  for (int k = 0; k < cnt; ++k) X->level[verts[k]] = -1;
  bfs(X, rid, root, queue, &nlev);
  int sepnode = -1;
  if (nlev >= 3) {
    /* level sizes; the separator level: near the median, narrowest among the levels holding the 40-60 % quantiles */
    int* lsize = (int*)calloc((size_t)nlev + 1, sizeof(int));
    for (int k = 0; k < cnt; ++k) lsize[X->level[verts[k]]]++;
    int best = -1, cum = 0;
    double best_cost = 1e300;
    for (int l = 0; l < nlev; ++l) {
      const int below = cum, above = cnt - cum - lsize[l];
      cum += lsize[l];
      if (l == 0 || l == nlev - 1 || below == 0 || above == 0) continue;
      const double imb = fabs((double)below - (double)above) / (double)cnt;   /* 0 = balanced */
      const double cost = (double)lsize[l] * (1.0 + 4.0 * imb * imb) + (imb > 0.35 ? 1e9 * imb : 0.0);
      if (cost < best_cost) { best_cost = cost; best = l; }
    }
    free(lsize);
    if (best > 0) {
      const int *rp = C->rowptr, *ci = C->colidx;
      /* S = vertices of level `best` with a neighbour in level best+1; the rest of the level joins the lower part */
      int *sep = (int*)malloc(sizeof(int) * (size_t)cnt), *lo = (int*)malloc(sizeof(int) * (size_t)cnt),
          *hi = (int*)malloc(sizeof(int) * (size_t)cnt);
      int ns = 0, nl = 0, nh = 0;
      for (int k = 0; k < cnt; ++k) {
        const int v = verts[k], lv = X->level[v];
        if (lv < best) lo[nl++] = v;
        else if (lv > best) hi[nh++] = v;
        else {
          int touches = 0;
          for (int q = rp[v]; q < rp[v + 1] && !touches; ++q) {
            const int u = ci[q];
            if (u != v && X->region[u] == rid && X->level[u] == best + 1) touches = 1;
          }
          if (touches) sep[ns++] = v; else lo[nl++] = v;
        }
      }
      if (ns > 0 && nl > 0 && nh > 0) {
        const int rlo = new_region(X), rhi = new_region(X);
        for (int k = 0; k < ns; ++k) X->region[sep[k]] = -1;
        for (int k = 0; k < nl; ++k) X->region[lo[k]] = rlo;
        for (int k = 0; k < nh; ++k) X->region[hi[k]] = rhi;
        ilist kids_lo = {0, 0, 0}, kids_hi = {0, 0, 0};
        if (cnt > 20000) {
#pragma omp task shared(kids_lo) firstprivate(lo, nl, rlo)
          nd_region(X, lo, nl, rlo, &kids_lo);
          nd_region(X, hi, nh, rhi, &kids_hi);
#pragma omp taskwait
        } else {
          nd_region(X, lo, nl, rlo, &kids_lo);
          nd_region(X, hi, nh, rhi, &kids_hi);
        }
        sepnode = new_node(C, ns, sep);
        const int nk = kids_lo.n + kids_hi.n;
        int* kids = (int*)malloc(sizeof(int) * (size_t)(nk > 0 ? nk : 1));
        memcpy(kids, kids_lo.ids, sizeof(int) * (size_t)kids_lo.n);
        memcpy(kids + kids_lo.n, kids_hi.ids, sizeof(int) * (size_t)kids_hi.n);
#pragma omp critical(chol_nodes)
        {
          C->nodes[sepnode].nchild = nk;
          C->nodes[sepnode].rel = kids;   /* borrowed until the child lists are flattened */
          for (int k = 0; k < nk; ++k) C->nodes[kids[k]].parent = sepnode;
        }
        free(kids_lo.ids);
        free(kids_hi.ids);
      }
      free(sep); free(lo); free(hi);
    }
  }
  if (sepnode < 0) {   /* no usable separator (clique-like component): one dense front */
    for (int k = 0; k < cnt; ++k) X->region[verts[k]] = -1;
    sepnode = new_node(C, cnt, verts);
  }
  free(queue);
  return sepnode;
}

/* a region may fall into several connected components: each heads its own subtree */
static void nd_region(nd_ctx* X, int* verts, int cnt, int rid, ilist* out) {
  if (cnt == 0) return;
  int* queue = (int*)malloc(sizeof(int) * (size_t)cnt);
  for (int k = 0; k < cnt; ++k) X->level[verts[k]] = -1;
  int first = 1;
  for (int k = 0; k < cnt; ++k) {
    const int v = verts[k];
    if (X->level[v] >= 0) continue;
    int nlev;
    const int m = bfs(X, rid, v, queue, &nlev);
    if (first && m == cnt) {   /* connected: the common case */
      ilist_push(out, nd_component(X, verts, cnt, rid));
      free(queue);
      return;
    }
    first = 0;
    const int sub = new_region(X);
    int* comp = (int*)malloc(sizeof(int) * (size_t)m);
    memcpy(comp, queue, sizeof(int) * (size_t)m);
    for (int q = 0; q < m; ++q) X->region[comp[q]] = sub;
    ilist_push(out, nd_component(X, comp, m, sub));
    for (int q = 0; q < m; ++q) X->level[comp[q]] = 0;   /* visited, for the rest of this scan */
    free(comp);
  }
  free(queue);
}

/* ------------------------------------------------------------------------------------------- symbolic */
static int cmp_int(const void* a, const void* b) {
  const int x = *(const int*)a, y = *(const int*)b;
  return (x > y) - (x < y);
}

static void postorder(chol* C) {
  /* flatten child lists, iterative DFS from the roots */
  C->child_list = (int*)malloc(sizeof(int) * (size_t)(C->nnodes > 0 ? C->nnodes : 1));
  int w = 0;
  for (int t = 0; t < C->nnodes; ++t) {
    node_t* nd = &C->nodes[t];
    nd->child0 = w;
    if (nd->nchild) {
      memcpy(C->child_list + w, nd->rel, sizeof(int) * (size_t)nd->nchild);
      free(nd->rel);
      nd->rel = 0;
    }
    w += nd->nchild;
  }
  C->post = (int*)malloc(sizeof(int) * (size_t)C->nnodes);
  int* stack = (int*)malloc(sizeof(int) * 2 * (size_t)C->nnodes);
  int np = 0;
  for (int r = 0; r < C->nroots; ++r) {
    int sp = 0;
    stack[sp++] = C->roots[r];
    stack[sp++] = 0;
    while (sp) {
      const int k = stack[sp - 1], t = stack[sp - 2];
      if (k < C->nodes[t].nchild) {
        stack[sp - 1] = k + 1;
        stack[sp++] = C->child_list[C->nodes[t].child0 + k];
        stack[sp++] = 0;
      } else {
        C->post[np++] = t;
        sp -= 2;
      }
    }
  }
  free(stack);
}

static int front_row(const node_t* p, int pos) {   /* row block of elimination position `pos` in p's front */
  if (pos < p->v0 + p->nv) return pos - p->v0;
  int lo = 0, hi = p->nb;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (p->bnd[mid] < pos) lo = mid + 1; else hi = mid;
  }
  return p->nv + lo;
}

chol* chol_analyze(int n, const int* rowptr, const int* colidx) {
  chol* C = (chol*)calloc(1, sizeof(chol));
  C->n = n;
  C->rowptr = rowptr;
  C->colidx = colidx;
  nd_ctx X;
  X.C = C;
  X.level = (int*)malloc(sizeof(int) * (size_t)n);
  X.region = (int*)calloc((size_t)n, sizeof(int));
  X.next_region = 1;
  int* all = (int*)malloc(sizeof(int) * (size_t)n);
  for (int v = 0; v < n; ++v) all[v] = v;
  ilist roots = {0, 0, 0};
#pragma omp parallel
#pragma omp single
  nd_region(&X, all, n, 0, &roots);
  free(all);
  free(X.level);
  free(X.region);
  C->roots = roots.ids;
  C->nroots = roots.n;
  postorder(C);
  /* elimination positions in postorder */
  C->perm = (int*)malloc(sizeof(int) * (size_t)n);
  C->iperm = (int*)malloc(sizeof(int) * (size_t)n);
  int pos = 0;
  for (int q = 0; q < C->nnodes; ++q) {
    node_t* nd = &C->nodes[C->post[q]];
    nd->v0 = pos;
    for (int k = 0; k < nd->nv; ++k) {
      C->perm[pos] = nd->verts[k];
      C->iperm[nd->verts[k]] = pos++;
    }
    free(nd->verts);
    nd->verts = 0;
  }
  if (pos != n) { C->failed = 1; return C; }
  /* boundaries, children before parents */
  for (int q = 0; q < C->nnodes; ++q) {
    node_t* nd = &C->nodes[C->post[q]];
    const int end = nd->v0 + nd->nv;
    int cap = 64, m = 0;
    int* tmp = (int*)malloc(sizeof(int) * (size_t)cap);
    for (int k = 0; k < nd->nv; ++k) {
      const int v = C->perm[nd->v0 + k];
      for (int e = rowptr[v]; e < rowptr[v + 1]; ++e) {
        const int p = C->iperm[colidx[e]];
        if (p < end) continue;
        if (m == cap) { cap *= 2; tmp = (int*)realloc(tmp, sizeof(int) * (size_t)cap); }
        tmp[m++] = p;
      }
    }
    for (int c = 0; c < nd->nchild; ++c) {
      const node_t* ch = &C->nodes[C->child_list[nd->child0 + c]];
      for (int k = 0; k < ch->nb; ++k) {
        if (ch->bnd[k] < end) continue;
        if (m == cap) { cap *= 2; tmp = (int*)realloc(tmp, sizeof(int) * (size_t)cap); }
        tmp[m++] = ch->bnd[k];
      }
    }
    qsort(tmp, (size_t)m, sizeof(int), cmp_int);
    int u = 0;
    for (int k = 0; k < m; ++k)
      if (u == 0 || tmp[k] != tmp[u - 1]) tmp[u++] = tmp[k];
    nd->nb = u;
    nd->bnd = tmp;
    const double f = 3.0 * (nd->nv + nd->nb), p3 = 3.0 * nd->nv;
    C->nnz_l += f * p3 - p3 * (p3 - 1) / 2;
    C->flops += p3 * f * f - p3 * p3 * f + p3 * p3 * p3 / 3.0;   /* sum_{k<p} (f-k)^2 */
    if (nd->nv + nd->nb > C->max_front) C->max_front = nd->nv + nd->nb;
  }
  for (int t = 0; t < C->nnodes; ++t) {
    node_t* nd = &C->nodes[t];
    nd->rel = (int*)malloc(sizeof(int) * (size_t)(nd->nb > 0 ? nd->nb : 1));
    if (nd->parent >= 0) {
      const node_t* p = &C->nodes[nd->parent];
      for (int k = 0; k < nd->nb; ++k) nd->rel[k] = front_row(p, nd->bnd[k]);
    }
    const size_t m3 = 3 * (size_t)(nd->nv + nd->nb);
    nd->L = (double*)malloc(sizeof(double) * m3 * 3 * (size_t)nd->nv);
    nd->y = (double*)malloc(sizeof(double) * m3);
  }
  return C;
}

/* ------------------------------------------------------------------------------------------- dense kernels */
/* C[0:mr, 0:nr] -= A[0:mr, 0:K] * B[0:nr, 0:K]^T, column-major with leading dimensions lda, ldb, ldc */
static inline void micro_full(int K, const double* restrict A, int lda, const double* restrict B, int ldb, double* restrict Cm,
                              int ldc) {
  double acc[NR][MR];
  for (int j = 0; j < NR; ++j)
    for (int i = 0; i < MR; ++i) acc[j][i] = 0.0;
  for (int k = 0; k < K; ++k) {
    const double* a = A + (size_t)k * lda;
    const double* b = B + (size_t)k * ldb;
    for (int j = 0; j < NR; ++j) {
      const double bj = b[j];
#pragma omp simd
      for (int i = 0; i < MR; ++i) acc[j][i] += a[i] * bj;
    }
  }
  for (int j = 0; j < NR; ++j) {
    double* c = Cm + (size_t)j * ldc;
#pragma omp simd
    for (int i = 0; i < MR; ++i) c[i] -= acc[j][i];
  }
}

static inline void micro_edge(int mr, int nr, int K, const double* A, int lda, const double* B, int ldb, double* Cm, int ldc) {
  for (int j = 0; j < nr; ++j)
    for (int i = 0; i < mr; ++i) {
      double s = 0.0;
      for (int k = 0; k < K; ++k) s += A[i + (size_t)k * lda] * B[j + (size_t)k * ldb];
      Cm[i + (size_t)j * ldc] -= s;
    }
}

/* lower triangle of C (r x r, ld) -= P P^T with P = r x K (ld): column tiles [j0, j1) */
static void syrk_tiles(int r, int K, const double* P, int ld, double* Cm, int j0, int j1) {
  for (int j = j0; j < j1; j += NR) {
    const int nr = j1 - j < NR ? j1 - j : NR;
    int i = j - j % MR;   /* first row tile that reaches the diagonal */
    if (i < 0) i = 0;
    for (; i < r; i += MR) {
      const int mr = r - i < MR ? r - i : MR;
      if (mr == MR && nr == NR) micro_full(K, P + i, ld, P + j, ld, Cm + i + (size_t)j * ld, ld);
      else micro_edge(mr, nr, K, P + i, ld, P + j, ld, Cm + i + (size_t)j * ld, ld);
    }
  }
}

/* in place: leading p columns of the m x m column-major lower-triangular F become [L11; L21], the trailing
 * (m-p) x (m-p) block its Schur complement.  Returns 0, or 1 + column of a non-positive pivot. */
static int partial_cholesky(double* F, int m, int p) {
  for (int k0 = 0; k0 < p; k0 += NB) {
    const int kb = p - k0 < NB ? p - k0 : NB;
    /* diagonal block: unblocked left-looking */
    for (int j = k0; j < k0 + kb; ++j) {
      double* cj = F + (size_t)j * m;
      for (int l = k0; l < j; ++l) {
        const double* cl = F + (size_t)l * m;
        const double f = cl[j];
        if (f != 0.0)
          for (int i = j; i < k0 + kb; ++i) cj[i] -= cl[i] * f;
      }
      const double d = cj[j];
      if (!(d > 0.0) || !isfinite(d)) return 1 + j;
      const double s = sqrt(d), inv = 1.0 / s;
      cj[j] = s;
      for (int i = j + 1; i < k0 + kb; ++i) cj[i] *= inv;
    }
    const int r0 = k0 + kb, r = m - r0;
    if (r == 0) break;
    /* panel below: X L11^T = F  ->  column by column (long contiguous axpys) */
    for (int j = k0; j < k0 + kb; ++j) {
      double* cj = F + (size_t)j * m + r0;
      for (int l = k0; l < j; ++l) {
        const double f = F[(size_t)l * m + j];
        const double* cl = F + (size_t)l * m + r0;
        if (f != 0.0) {
#pragma omp simd
          for (int i = 0; i < r; ++i) cj[i] -= cl[i] * f;
        }
      }
      const double inv = 1.0 / F[(size_t)j * m + j];
#pragma omp simd
      for (int i = 0; i < r; ++i) cj[i] *= inv;
    }
    /* trailing update: lower(F[r0:, r0:]) -= P P^T */
    const double* P = F + (size_t)k0 * m + r0;
    double* T = F + (size_t)r0 * m + r0;
    if ((double)r * r * kb > 4e7) {
      const int step = NR * 16;
#pragma omp taskloop grainsize(1) default(shared) firstprivate(r, kb, P, T, m, step)
      for (int j = 0; j < r; j += step) syrk_tiles(r, kb, P, m, T, j, j + step < r ? j + step : r);
    } else {
      syrk_tiles(r, kb, P, m, T, 0, r);
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------- numeric */
typedef struct {
  chol* C;
  const double* H;       /* 9 per block of the full BSR (row-major blocks), pattern rowptr/colidx */
  const int* diag;       /* per row: index of its diagonal block */
  double lambda;
  const double* dscale;  /* NULL: lambda*I, else lambda*diag(dscale) (3 per vertex) */
  int fail;
} num_ctx;

static void factor_node(num_ctx* N, int t) {
  chol* C = N->C;
  node_t* nd = &C->nodes[t];
  const int nf = nd->nv + nd->nb, m = 3 * nf, p = 3 * nd->nv;
  double* F = (double*)calloc((size_t)m * m, sizeof(double));
  if (!F) { N->fail = 1; return; }
  /* original entries of the pivot columns */
  for (int k = 0; k < nd->nv; ++k) {
    const int v = C->perm[nd->v0 + k];
    for (int e = C->rowptr[v]; e < C->rowptr[v + 1]; ++e) {
      const int u = C->colidx[e];
      const int pu = C->iperm[u];
      if (pu < nd->v0 + k) continue;                       /* upper part / eliminated earlier */
      const int ri = front_row(nd, pu);
      const double* h = N->H + 9 * (size_t)e;              /* H_vu, row-major; entry (u,v) of the lower triangle = H_vu^T */
      for (int a = 0; a < 3; ++a)                           /* a: component of u (row), b: component of v (column) */
        for (int b = 0; b < 3; ++b) {
          if (u == v && a < b) continue;
          F[(size_t)(3 * ri + a) + (size_t)(3 * k + b) * m] = h[3 * b + a];
        }
      if (u == v)
        for (int a = 0; a < 3; ++a)
          F[(size_t)(3 * k + a) * (m + 1)] += N->lambda * (N->dscale ? N->dscale[3 * (size_t)v + a] : 1.0);
    }
  }
  /* extend-add the children's update matrices */
  for (int c = 0; c < nd->nchild; ++c) {
    node_t* ch = &C->nodes[C->child_list[nd->child0 + c]];
    const int mb = 3 * ch->nb;
    if (!ch->U) continue;
    for (int j = 0; j < ch->nb; ++j) {
      const int cj = ch->rel[j];
      for (int i = j; i < ch->nb; ++i) {
        const int ri = ch->rel[i];   /* rel ascending with i, so ri >= cj: lower triangle */
        for (int b = 0; b < 3; ++b)
          for (int a = 0; a < 3; ++a) {
            if (i == j && a < b) continue;
            F[(size_t)(3 * ri + a) + (size_t)(3 * cj + b) * m] += ch->U[(size_t)(3 * i + a) + (size_t)(3 * j + b) * mb];
          }
      }
    }
    free(ch->U);
    ch->U = 0;
  }
  const int bad = partial_cholesky(F, m, p);
  if (bad) N->fail = 1;
  /* keep [L11; L21]; hand the Schur complement to the parent */
  memcpy(nd->L, F, sizeof(double) * (size_t)m * p);
  if (nd->nb && nd->parent >= 0 && !bad) {
    const int mb = 3 * nd->nb;
    nd->U = (double*)malloc(sizeof(double) * (size_t)mb * mb);
    for (int j = 0; j < mb; ++j) memcpy(nd->U + (size_t)j * mb + j, F + (size_t)(p + j) * m + p + j, sizeof(double) * (size_t)(mb - j));
  }
  free(F);
}

static void factor_subtree(num_ctx* N, int t) {
  chol* C = N->C;
  node_t* nd = &C->nodes[t];
  for (int c = 0; c < nd->nchild; ++c) {
    const int ch = C->child_list[nd->child0 + c];
    const node_t* cn = &C->nodes[ch];
    if (cn->nv + cn->nb >= 24 && cn->nchild) {
#pragma omp task default(shared) firstprivate(ch)
      factor_subtree(N, ch);
    } else {
      factor_subtree(N, ch);
    }
  }
#pragma omp taskwait
  if (!N->fail) factor_node(N, t);
}

int chol_factor(chol* C, const double* H, const int* diag, double lambda, const double* dscale) {
  if (C->failed) return 1;
  num_ctx N;
  N.C = C; N.H = H; N.diag = diag; N.lambda = lambda; N.dscale = dscale; N.fail = 0;
#pragma omp parallel
#pragma omp single
  for (int r = 0; r < C->nroots; ++r) {
#pragma omp task default(shared) firstprivate(r)
    factor_subtree(&N, C->roots[r]);
  }
  for (int t = 0; t < C->nnodes; ++t)
    if (C->nodes[t].U) { free(C->nodes[t].U); C->nodes[t].U = 0; }
  return N.fail;
}

/* ------------------------------------------------------------------------------------------- solve */
static void fwd_subtree(chol* C, int t, const double* rhs) {
  node_t* nd = &C->nodes[t];
  for (int c = 0; c < nd->nchild; ++c) {
    const int ch = C->child_list[nd->child0 + c];
    if (C->nodes[ch].nchild) {
#pragma omp task default(shared) firstprivate(ch)
      fwd_subtree(C, ch, rhs);
    } else {
      fwd_subtree(C, ch, rhs);
    }
  }
#pragma omp taskwait
  const int m = 3 * (nd->nv + nd->nb), p = 3 * nd->nv;
  double* y = nd->y;
  for (int k = 0; k < nd->nv; ++k) {
    const int v = C->perm[nd->v0 + k];
    y[3 * k] = rhs[3 * (size_t)v]; y[3 * k + 1] = rhs[3 * (size_t)v + 1]; y[3 * k + 2] = rhs[3 * (size_t)v + 2];
  }
  for (int i = p; i < m; ++i) y[i] = 0.0;
  for (int c = 0; c < nd->nchild; ++c) {   /* children's updates, in child order (deterministic) */
    const node_t* ch = &C->nodes[C->child_list[nd->child0 + c]];
    const double* u = ch->y + 3 * ch->nv;
    for (int i = 0; i < ch->nb; ++i) {
      const int ri = ch->rel[i];
      y[3 * ri] += u[3 * i]; y[3 * ri + 1] += u[3 * i + 1]; y[3 * ri + 2] += u[3 * i + 2];
    }
  }
  /* y_V = L11^-1 y_V ; y_B -= L21 y_V  (column sweep) */
  for (int j = 0; j < p; ++j) {
    const double* lj = nd->L + (size_t)j * m;
    const double yj = y[j] / lj[j];
    y[j] = yj;
#pragma omp simd
    for (int i = j + 1; i < m; ++i) y[i] -= lj[i] * yj;
  }
}

static void bwd_subtree(chol* C, int t, double* x) {
  node_t* nd = &C->nodes[t];
  const int m = 3 * (nd->nv + nd->nb), p = 3 * nd->nv;
  double* y = nd->y;
  /* boundary values from the ancestors (already final in x, by elimination position) */
  for (int i = 0; i < nd->nb; ++i) {
    const int v = C->perm[nd->bnd[i]];
    y[p + 3 * i] = x[3 * (size_t)v]; y[p + 3 * i + 1] = x[3 * (size_t)v + 1]; y[p + 3 * i + 2] = x[3 * (size_t)v + 2];
  }
  for (int j = p - 1; j >= 0; --j) {
    const double* lj = nd->L + (size_t)j * m;
    double s = y[j];
#pragma omp simd reduction(- : s)
    for (int i = j + 1; i < m; ++i) s -= lj[i] * y[i];
    y[j] = s / lj[j];
  }
  for (int k = 0; k < nd->nv; ++k) {
    const int v = C->perm[nd->v0 + k];
    x[3 * (size_t)v] = y[3 * k]; x[3 * (size_t)v + 1] = y[3 * k + 1]; x[3 * (size_t)v + 2] = y[3 * k + 2];
  }
  for (int c = 0; c < nd->nchild; ++c) {
    const int ch = C->child_list[nd->child0 + c];
    if (C->nodes[ch].nchild) {
#pragma omp task default(shared) firstprivate(ch)
      bwd_subtree(C, ch, x);
    } else {
      bwd_subtree(C, ch, x);
    }
  }
#pragma omp taskwait
}

void chol_solve(chol* C, const double* rhs, double* x) {
#pragma omp parallel
#pragma omp single
  {
    for (int r = 0; r < C->nroots; ++r) {
#pragma omp task default(shared) firstprivate(r)
      fwd_subtree(C, C->roots[r], rhs);
    }
#pragma omp taskwait
    for (int r = 0; r < C->nroots; ++r) {
#pragma omp task default(shared) firstprivate(r)
      bwd_subtree(C, C->roots[r], x);
    }
  }
}

void chol_info(const chol* C, double* nnz_l, double* flops, int* max_front, int* nnodes) {
  if (nnz_l) *nnz_l = C->nnz_l;
  if (flops) *flops = C->flops;
  if (max_front) *max_front = C->max_front;
  if (nnodes) *nnodes = C->nnodes;
}

void chol_free(chol* C) {
  if (!C) return;
  for (int t = 0; t < C->nnodes; ++t) {
    node_t* nd = &C->nodes[t];
    free(nd->verts); free(nd->bnd); free(nd->rel); free(nd->L); free(nd->U); free(nd->y);
  }
  free(C->nodes); free(C->child_list); free(C->roots); free(C->post); free(C->perm); free(C->iperm);
  free(C);
}
