// Host-side structure build (K6): the one-off, integer-only work that GTSAM redoes on every
// LevenbergMarquardtOptimizer construction (Ordering / VariableIndex / symbolic factorisation;
// reference src/graph/src/graph.cpp:119).  Pure C++ (no CUDA), so the CPU tests exercise exactly
// the code the product runs.
//
// Everything is derived from one canonical, insertion-order-independent order of the factors,
//     sort by (min(i,j), max(i,j), direction, insertion index),
// which fixes every summation order downstream (no atomics anywhere, bit-reproducible):
//   * vertex-range partition : rows are cut into kSegments fixed segments on kTile boundaries; rank r
//                              of P owns kSegments/P consecutive segments.  Every reduction is
//                              tile -> segment -> total in that fixed tree, so results are bitwise
//                              identical for P = 1, 2, 4, 8.
//   * factor storage         : SoA; every contribution slot owns a cell, the factor's data sits in the
//                              cell of its "home" row = the smaller endpoint (the larger one when the
//                              smaller is owned by another rank); see "Cells" below.
//   * contribution lists     : per owned row, every incident factor side ordered by the other
//                              endpoint's global index, then canonical order; SELL-32 laid out so a
//                              warp of 32 consecutive rows reads 32 consecutive words per slot.
//   * H off-diagonal blocks  : ONE 3x3 block per unordered vertex pair, stored in canonical
//                              orientation M = A_lo^T Omega A_hi (lo = smaller global index) in the
//                              reference cell of its writer row; the other row uses M^T (shared or copied).
//   * reference lists        : per owned row, its neighbours in global column order as
//                              (local column, block position, transposed?) for the SpMV.
//   * halo lists             : for each peer, the owned rows it needs (send) and the contiguous
//                              range of local ghost slots it fills (recv).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "host_pool.hpp"

namespace gs {

constexpr int kTile = 256;      // rows per CTA = rows per reduction tile
constexpr int kSlice = 32;      // rows per SELL slice = one warp
constexpr int kSegments = 8;    // fixed reduction segments; world size must divide it
constexpr uint32_t kPadCode = 0xFFFFFFFFu;

// contribution code: (factor position << 3) | (count chi2 here << 2) | kind
//   kind 0: row is id_1 of the factor, 1: row is id_2, 2: prior (position = prior index)
// contribution block word: -1, or (block position << 2) | (row is the hi endpoint << 1) | head
// reference block word: (block position << 1) | transposed

struct Partition {
  int N = 0, P = 1;
  int seg_row[kSegments + 1] = {0};  // global row boundaries (multiples of kTile, last = N)
  void init(int n, int world) {
    N = n;
    P = world;
    const int64_t tiles = ((int64_t)n + kTile - 1) / kTile;
    for (int k = 0; k <= kSegments; ++k) seg_row[k] = (int)std::min<int64_t>(n, (tiles * k / kSegments) * kTile);
  }
  int segs_per_rank() const { return kSegments / P; }
  int row_begin(int rank) const { return seg_row[rank * segs_per_rank()]; }
  int row_end(int rank) const { return seg_row[(rank + 1) * segs_per_rank()]; }
  int owner(int row) const {
    for (int r = 0; r < P; ++r)
      if (row < row_end(r)) return r;
    return P - 1;
  }
};

// sorted unique 3x3 block pattern of H including the diagonal, over global dense indices
// (the integer parity hook; independent of the partition)
inline void global_pattern(int N, int E, const int* ei, const int* ej, std::vector<int32_t>& rowptr,
                           std::vector<int32_t>& colidx) {
  std::vector<int> cnt(N + 1, 0);
  for (int k = 0; k < E; ++k) {
    cnt[ei[k] + 1]++;
    cnt[ej[k] + 1]++;
  }
  for (int r = 0; r < N; ++r) cnt[r + 1] += cnt[r];
  std::vector<int> adj((size_t)2 * E), fill(cnt.begin(), cnt.end() - 1);
  for (int k = 0; k < E; ++k) {
    adj[fill[ei[k]]++] = ej[k];
    adj[fill[ej[k]]++] = ei[k];
  }
  rowptr.assign(N + 1, 0);
  colidx.clear();
  colidx.reserve((size_t)2 * E + N);
  for (int r = 0; r < N; ++r) {
    rowptr[r] = (int32_t)colidx.size();
    std::sort(adj.begin() + cnt[r], adj.begin() + cnt[r + 1]);
    bool diag_done = false;
    int last = -1;
    for (int q = cnt[r]; q < cnt[r + 1]; ++q) {
      const int c = adj[q];
      if (c == last) continue;
      if (!diag_done && c > r) {
        colidx.push_back(r);
        diag_done = true;
      }
      colidx.push_back(c);
      last = c;
    }
    if (!diag_done) colidx.push_back(r);
  }
  rowptr[N] = (int32_t)colidx.size();
}

// canonical factor order: k-th canonical factor = caller's factor perm[k]
inline void canonical_order(int N, int E, const int* ei, const int* ej, std::vector<int>& perm) {
  std::vector<int> gptr(N + 1, 0);
  for (int k = 0; k < E; ++k) gptr[std::min(ei[k], ej[k]) + 1]++;
  for (int r = 0; r < N; ++r) gptr[r + 1] += gptr[r];
  perm.resize(E);
  {
    std::vector<int> fill(gptr.begin(), gptr.end() - 1);
    for (int k = 0; k < E; ++k) perm[fill[std::min(ei[k], ej[k])]++] = k;
  }
  for (int r = 0; r < N; ++r) {
    const int a = gptr[r], b = gptr[r + 1];
    if (b - a > 1) {
      std::sort(perm.begin() + a, perm.begin() + b, [&](int x, int y) {
        const int mx = std::max(ei[x], ej[x]), my = std::max(ei[y], ej[y]);
        if (mx != my) return mx < my;
        const int dx = ei[x] > ej[x], dy = ei[y] > ej[y];
        if (dx != dy) return dx < dy;
        return x < y;
      });
    }
  }
}

// SELL-32 layout over per-row counts: slot k of row r lives at base[r/32] + k*32 + r%32
struct Sell {
  std::vector<int> base, width;  // per slice
  int slots = 0;
  void build(const std::vector<int>& cnt) {
    const int n = (int)cnt.size();
    const int ns = (n + kSlice - 1) / kSlice;
    base.assign(ns + 1, 0);
    width.assign(std::max(ns, 1), 0);
    int64_t tot = 0;
    for (int s = 0; s < ns; ++s) {
      int w = 0;
      for (int r = s * kSlice; r < std::min(n, (s + 1) * kSlice); ++r) w = std::max(w, cnt[r]);
      width[s] = w;
      base[s] = (int)tot;
      tot += (int64_t)w * kSlice;
    }
    base[ns] = (int)tot;
    slots = (int)tot;
  }
  int pos(int row, int k) const { return base[row / kSlice] + k * kSlice + (row % kSlice); }
};

struct HaloPeer {
  int rank = 0;
  int send_off = 0, send_cnt = 0;   // range in send_rows
  int ghost_off = 0, ghost_cnt = 0; // range of ghost slots (local index = n_own + ghost_off + k)
};

// Cells.  Every (row, slot) of a SELL list owns a storage cell at the same position: contribution
// (row, s) owns factor cell csell.pos(row, s), reference (row, s) owns block cell rsell.pos(row, s).
// A cell either HOLDS data (the factor / block is "homed" there, or copied there) or POINTS at the
// home cell of another row.  Pointing shares bytes (one DRAM read serves both rows, the second comes
// from L2) but is only chosen when the 32 rows of a slice point at nearby cells, i.e. when the
// warp's gather stays coalesced; otherwise the structure build gives the cell its own copy, so
// every warp-level access of every kernel touches a handful of 128-byte lines whatever the graph.
constexpr int kShareMaxLinesFactor = 12;  // distinct 128-B lines (8 double2 per line) a slice-slot may gather factors from
constexpr int kShareMaxLinesBlock = 8;    // distinct 128-B lines (16 doubles per line) a slice-slot may gather blocks from

struct LocalStructure {
  Partition part;
  int rank = 0;
  int v0 = 0, v1 = 0;        // owned global rows [v0, v1)
  int n_own = 0, n_ghost = 0;
  int E_glob = 0, E_loc = 0, P_loc = 0;
  int n_pairs = 0;           // off-diagonal blocks computed here (unordered pairs with an owned endpoint)
  int nnzo_own = 0;          // off-diagonal blocks of the owned rows of the full (non-symmetric) BSR
  int n_factor_copies = 0;   // factor cells that hold a copy (gather would not coalesce)
  int n_block_copies = 0;    // block cells that hold a transposed copy
  int n_merged = 0;          // block contributions that continue a run (a second factor between the same pair); 0 =
                             // every stored block comes from exactly one factor
  int ntiles = 0;
  int seg_tile[kSegments + 1] = {0};  // local tile ranges of the locally owned segments (first kSegments/P used)
  std::vector<int> ghosts;            // global index of ghost g (sorted)

  // contribution lists = factor cells
  Sell csell;
  std::vector<uint32_t> c_code;       // (factor cell to read << 3) | count << 2 | kind ; kPadCode padding
  std::vector<int> c_blk;             // -1, or (block cell << 2) | (row is hi endpoint << 1) | head
  std::vector<int> c_first;           // per owned row: first slot whose cell is a counting side (col > row)
  std::vector<int> c_col;             // local index of the other vertex (0 for priors/padding)
  std::vector<int> f_src;             // factor cell -> caller's factor index, -1 if the cell holds no data
  std::vector<int> f_i, f_j;          // local vertex index of id_1 / id_2 per data cell (-1 otherwise)
  std::vector<char> f_count;          // 1: this cell is the counting side of its factor (chi2)
  // priors on owned rows
  std::vector<int> pr_src;            // local prior -> caller's prior index
  std::vector<int> pr_row;            // local row, ascending
  // reference lists = block cells
  Sell rsell;
  std::vector<int> r_col, r_blk;      // local column (-1 padding); (block cell to read << 1) | transposed
  // halo
  std::vector<HaloPeer> peers;
  std::vector<int> send_rows;         // local owned row indices, grouped by peer, each group sorted
  int max_degree = 0;
  // scratch of build_local_structure, kept between builds: an incremental update rebuilds the structure every few
  // keyframes, and fresh vectors of this size cost a page fault per 4 KB each time
  struct Work {
    std::vector<int> loc, ccnt, ccol, chome, cfac, fill, rcnt, rptr, nbr, rhome;
  } work;

  int local_of(int g) const {
    if (g >= v0 && g < v1) return g - v0;
    const auto it = std::lower_bound(ghosts.begin(), ghosts.end(), g);
    return n_own + (int)(it - ghosts.begin());
  }
};

// ---- coarse space of the two-level preconditioner (host side, integer-only) ---------------------------------------
// The block rows are grouped into AGGREGATES of about `target` graph-adjacent vertices; each aggregate carries three
// coarse unknowns (its rigid-body motion).  Aggregates never cross the fixed segment boundaries of the partition, so an
// aggregate belongs to one rank whatever the number of ranks, and the aggregation does not depend on that number.
// Built by recursive bisection: a breadth-first order of the vertex set from a far-away vertex (two sweeps), cut at
// its median, until a set has at most `target` vertices (two sweeps per cut: one to find the far-away vertex, one for the order).  Deterministic (no hashing, no randomness).
//   colour : distance-2 colouring of the aggregate adjacency graph — aggregates of one colour have no common
//            neighbour, so the coarse matrix E = W^T H W is formed with one product H w per (colour, rigid mode)
//            instead of one per aggregate.
struct Aggregation {
  int n_agg = 0, n_colors = 0, target = 0;
  std::vector<int> agg_of;        // per global row
  std::vector<int> ptr, rows;     // rows of aggregate a: rows[ptr[a] .. ptr[a+1]), ascending; rows[ptr[a]] is its reference row
  std::vector<int> color;         // per aggregate
  std::vector<int> nbr_of_color;  // [a * n_colors + c]: the aggregate of colour c in the closed neighbourhood of a, or -1
  int seg_agg[kSegments + 1] = {0};
  std::vector<int> work_region, work_mark;   // scratch of build_aggregation, kept between builds (no fresh pages per build)
};

// One segment's aggregates (each a sorted row list), in the order the bisection produces them.  Touches only the
// entries [r0, r1) of region / mark, so the eight segments can run on eight host threads.
inline void aggregate_segment(int r0, int r1, const std::vector<int32_t>& rowptr, const std::vector<int32_t>& colidx, int target,
                              std::vector<int>& region, std::vector<int>& mark, std::vector<std::vector<int>>& out) {
  out.clear();
  if (r1 <= r0) return;
  std::vector<int> order;
  int next_region = 1, stamp = 0;
  // breadth-first order of `verts` (all of region rid), component after component, starting at `root`
  auto bfs_order = [&](const std::vector<int>& verts, int rid, int root, std::vector<int>& o) {
    ++stamp;
    o.clear();
    size_t scan = 0;
    int start = root;
    for (;;) {
      size_t head = o.size();
      o.push_back(start);
      mark[start] = stamp;
      while (head < o.size()) {
        const int v = o[head++];
        for (int k = rowptr[v]; k < rowptr[v + 1]; ++k) {
          const int u = colidx[k];
          if (u < r0 || u >= r1 || region[u] != rid || mark[u] == stamp) continue;
          mark[u] = stamp;
          o.push_back(u);
        }
      }
      if (o.size() == verts.size()) break;
      while (mark[verts[scan]] == stamp) ++scan;   // next component
      start = verts[scan];
    }
  };
  struct Job {
    std::vector<int> verts;
    int rid;
  };
  std::vector<Job> stack;
  {
    Job j;
    j.rid = next_region++;
    j.verts.resize(r1 - r0);
    for (int r = r0; r < r1; ++r) {
      j.verts[r - r0] = r;
      region[r] = j.rid;
      mark[r] = 0;
    }
    stack.push_back(std::move(j));
  }
  while (!stack.empty()) {
    Job j = std::move(stack.back());
    stack.pop_back();
    const int cnt = (int)j.verts.size();
    if (cnt <= target) {
      std::sort(j.verts.begin(), j.verts.end());
      for (int v : j.verts) region[v] = -1;
      out.push_back(std::move(j.verts));
      continue;
    }
    bfs_order(j.verts, j.rid, j.verts[0], order);
    const int far = order.back();                   // a vertex far from an arbitrary one: (pseudo-)peripheral
    bfs_order(j.verts, j.rid, far, order);
    const int half = cnt / 2;
    Job lo, hi;
    lo.rid = next_region++;
    hi.rid = next_region++;
    lo.verts.assign(order.begin(), order.begin() + half);
    hi.verts.assign(order.begin() + half, order.end());
    for (int v : lo.verts) region[v] = lo.rid;
    for (int v : hi.verts) region[v] = hi.rid;
    stack.push_back(std::move(hi));   // lo is processed first
    stack.push_back(std::move(lo));
  }
}

inline void build_aggregation(int N, const std::vector<int32_t>& rowptr, const std::vector<int32_t>& colidx, int target,
                              Aggregation& A) {
  Partition part;
  part.init(N, 1);
  A.target = target;
  A.n_agg = 0;
  A.agg_of.assign(N, -1);
  A.ptr.assign(1, 0);
  A.rows.clear();
  A.rows.reserve(N);
  A.work_region.resize(N);
  A.work_mark.resize(N);
  std::vector<std::vector<int>> seg_out[kSegments];
  {
    // the segments are independent: one host thread each on graphs where a segment is worth a thread
    auto work = [&](int sgm) {
      aggregate_segment(part.seg_row[sgm], part.seg_row[sgm + 1], rowptr, colidx, target, A.work_region, A.work_mark, seg_out[sgm]);
    };
    if (N >= 8192) HostPool::get().run(kSegments, work);
    else
      for (int sgm = 0; sgm < kSegments; ++sgm) work(sgm);
  }
  for (int sgm = 0; sgm < kSegments; ++sgm) {
    A.seg_agg[sgm] = A.n_agg;
    for (const std::vector<int>& verts : seg_out[sgm]) {
      for (int v : verts) {
        A.agg_of[v] = A.n_agg;
        A.rows.push_back(v);
      }
      A.ptr.push_back((int)A.rows.size());
      A.n_agg++;
    }
  }
  A.seg_agg[kSegments] = A.n_agg;
  // aggregate adjacency (closed neighbourhoods), then greedy distance-2 colouring in aggregate order
  const int na = A.n_agg;
  std::vector<std::vector<int>> adj(na);
  {
    auto adjacency = [&](int a_begin, int a_end) {
      for (int a = a_begin; a < a_end; ++a) {
        std::vector<int>& l = adj[a];
        int last = -1;
        for (int q = A.ptr[a]; q < A.ptr[a + 1]; ++q) {
          const int v = A.rows[q];
          for (int k = rowptr[v]; k < rowptr[v + 1]; ++k) {
            const int b = A.agg_of[colidx[k]];
            if (b != last) l.push_back(b);   // most neighbours lie in the row's own aggregate
            last = b;
          }
        }
        std::sort(l.begin(), l.end());
        l.erase(std::unique(l.begin(), l.end()), l.end());   // contains a itself (diagonal entry)
      }
    };
    const int nt = N >= 8192 ? std::min(kSegments, na) : 1;
    HostPool::get().run(nt, [&](int t) { adjacency((int)((int64_t)na * t / nt), (int)((int64_t)na * (t + 1) / nt)); });
  }
  A.color.assign(na, -1);
  A.n_colors = 0;
  std::vector<int> forbid;
  for (int a = 0; a < na; ++a) {
    forbid.clear();
    for (int b : adj[a])
      for (int c2 : adj[b])
        if (A.color[c2] >= 0) forbid.push_back(A.color[c2]);
    std::sort(forbid.begin(), forbid.end());
    int c = 0;
    for (int f : forbid) {
      if (f == c) ++c;
      else if (f > c) break;
    }
    A.color[a] = c;
    if (c + 1 > A.n_colors) A.n_colors = c + 1;
  }
  A.nbr_of_color.assign((size_t)na * A.n_colors, -1);
  for (int a = 0; a < na; ++a)
    for (int b : adj[a]) A.nbr_of_color[(size_t)a * A.n_colors + A.color[b]] = b;
}

// fn(begin, end, part) over [0, n) in up to kSegments contiguous parts of at least `grain` items, one host thread each;
// returns the number of parts (sums a pass needs are kept per part and added in part order)
template <typename F>
inline int for_parts(int n, int grain, F fn) {
  const int parts = std::max(1, std::min(kSegments, n / std::max(grain, 1)));
  if (parts == 1) {
    fn(0, n, 0);
    return 1;
  }
  HostPool::get().run(parts, [&](int t) { fn((int)((int64_t)n * t / parts), (int)((int64_t)n * (t + 1) / parts), t); });
  return parts;
}
constexpr int kRowGrain = 8192;     // rows per host thread in the per-row passes of the structure build
constexpr int kSliceGrain = kRowGrain / kSlice;

// distinct values of key(x) over a small array (<= 32 entries)
template <typename F>
inline int distinct_count(const int* v, int n, F key) {
  int seen[kSlice];
  int m = 0;
  for (int a = 0; a < n; ++a) {
    const int k = key(v[a]);
    bool dup = false;
    for (int b = 0; b < m; ++b)
      if (seen[b] == k) {
        dup = true;
        break;
      }
    if (!dup) seen[m++] = k;
  }
  return m;
}

// Build the structure of `rank` in a world of `world` ranks.  ei/ej are global dense indices in the
// caller's factor order; `perm` is canonical_order() of them (shared by all ranks of a process).
inline void build_local_structure(int N, int E, const int* ei, const int* ej, const std::vector<int>& perm, int P,
                                  const int* prior_idx, int world, int rank, LocalStructure& S) {
  S.part.init(N, world);
  S.rank = rank;
  S.v0 = S.part.row_begin(rank);
  S.v1 = S.part.row_end(rank);
  S.n_own = S.v1 - S.v0;
  S.E_glob = E;
  const int v0 = S.v0, v1 = S.v1, n_own = S.n_own;
  auto owned = [&](int g) { return g >= v0 && g < v1; };

  // ---- local factors in canonical order, ghosts
  std::vector<int>& loc = S.work.loc;  // canonical positions (index into perm) of local factors
  loc.clear();
  loc.reserve(world == 1 ? (size_t)E : (size_t)E / world + 1024);
  S.ghosts.clear();
  for (int c = 0; c < E; ++c) {
    const int k = perm[c];
    const bool oi = owned(ei[k]), oj = owned(ej[k]);
    if (!oi && !oj) continue;
    loc.push_back(c);
    if (!oi) S.ghosts.push_back(ei[k]);
    if (!oj) S.ghosts.push_back(ej[k]);
  }
  std::sort(S.ghosts.begin(), S.ghosts.end());
  S.ghosts.erase(std::unique(S.ghosts.begin(), S.ghosts.end()), S.ghosts.end());
  S.n_ghost = (int)S.ghosts.size();
  S.E_loc = (int)loc.size();
  const int EL = S.E_loc;

  // ---- priors on owned rows: ascending row, caller's order within a row
  S.pr_src.clear();
  S.pr_row.clear();
  for (int p = 0; p < P; ++p)
    if (owned(prior_idx[p])) S.pr_src.push_back(p);
  std::stable_sort(S.pr_src.begin(), S.pr_src.end(), [&](int a, int b) { return prior_idx[a] < prior_idx[b]; });
  for (int p : S.pr_src) S.pr_row.push_back(prior_idx[p] - v0);
  S.P_loc = (int)S.pr_src.size();

  // ---- contribution lists: priors first, then factor sides in canonical order (which visits a
  //      row's neighbours in ascending global column order)
  std::vector<int>& ccnt = S.work.ccnt;
  ccnt.assign(n_own, 0);
  for (int p = 0; p < S.P_loc; ++p) ccnt[S.pr_row[p]]++;
  for (int q = 0; q < EL; ++q) {
    const int k = perm[loc[q]];
    if (owned(ei[k])) ccnt[ei[k] - v0]++;
    if (owned(ej[k])) ccnt[ej[k] - v0]++;
  }
  S.csell.build(ccnt);
  const int CS = S.csell.slots;
  S.c_code.assign(CS, kPadCode);
  S.c_blk.assign(CS, -1);
  S.c_col.assign(CS, 0);
  S.f_src.assign(CS, -1);
  S.f_i.assign(CS, -1);
  S.f_j.assign(CS, -1);
  S.f_count.assign(CS, 0);
  std::vector<int>&ccol = S.work.ccol, &chome = S.work.chome, &cfac = S.work.cfac;
  ccol.assign(CS, -1);    // global column of each contribution (-1 prior / padding)
  chome.assign(CS, -1);   // home cell of the contribution's factor
  cfac.assign(CS, -1);    // caller's factor index
  {
    std::vector<int>& fill = S.work.fill;
    fill.assign(n_own, 0);
    for (int p = 0; p < S.P_loc; ++p) {
      const int row = S.pr_row[p];
      const int c = S.csell.pos(row, fill[row]++);
      S.c_code[c] = ((uint32_t)p << 3) | 4u | 2u;
    }
    for (int q = 0; q < EL; ++q) {
      const int k = perm[loc[q]];
      const int gi = ei[k], gj = ej[k];
      const int lo = std::min(gi, gj), hi = std::max(gi, gj);
      const int ci = owned(gi) ? S.csell.pos(gi - v0, fill[gi - v0]++) : -1;
      const int cj = owned(gj) ? S.csell.pos(gj - v0, fill[gj - v0]++) : -1;
      const int home = owned(lo) ? (gi == lo ? ci : cj) : (gi == hi ? ci : cj);
      if (ci >= 0) {
        S.c_code[ci] = (gi == lo ? 4u : 0u) | 0u;
        ccol[ci] = gj;
        chome[ci] = home;
        cfac[ci] = k;
      }
      if (cj >= 0) {
        S.c_code[cj] = (gj == lo ? 4u : 0u) | 1u;
        ccol[cj] = gi;
        chome[cj] = home;
        cfac[cj] = k;
      }
    }
  }
  // share or copy, per (slice, slot): cells whose factor is homed elsewhere point there only if the
  // slice's 32 pointers stay within a few lines
  S.n_factor_copies = 0;
  {
    // a slice's cells are written by the part that owns the slice only (chome of its own cells), then finished there
    const int ns = (n_own + kSlice - 1) / kSlice;
    int sums[kSegments] = {0};
    for_parts(ns, kSliceGrain, [&](int sl0, int sl1, int part) {
      int away[kSlice], awaycell[kSlice];
      int copies = 0;
      for (int sl = sl0; sl < sl1; ++sl)
        for (int s = 0; s < S.csell.width[sl]; ++s) {
          int na = 0;
          for (int l = 0; l < kSlice && sl * kSlice + l < n_own; ++l) {
            const int c = S.csell.base[sl] + s * kSlice + l;
            if (cfac[c] >= 0 && chome[c] != c) {
              away[na] = chome[c];
              awaycell[na++] = c;
            }
          }
          if (na && distinct_count(away, na, [](int p) { return p >> 3; }) > kShareMaxLinesFactor) {
            for (int a = 0; a < na; ++a) chome[awaycell[a]] = awaycell[a];  // keep a copy in the own cell
            copies += na;
          }
        }
      for (int c = S.csell.base[sl0]; c < S.csell.base[sl1]; ++c) {
        if (cfac[c] < 0) continue;
        const int k = cfac[c];
        S.c_code[c] |= (uint32_t)chome[c] << 3;
        S.c_col[c] = S.local_of(ccol[c]);
        if (chome[c] == c) {  // data cell
          S.f_src[c] = k;
          S.f_i[c] = S.local_of(ei[k]);
          S.f_j[c] = S.local_of(ej[k]);
          S.f_count[c] = (S.c_code[c] & 4u) ? 1 : 0;
        }
      }
      sums[part] = copies;
    });
    for (int t = 0; t < kSegments; ++t) S.n_factor_copies += sums[t];
  }

  // ---- reference lists = unique neighbours per owned row, ascending global column (the
  //      contribution lists are already in that order)
  std::vector<int>& rcnt = S.work.rcnt;
  rcnt.assign(n_own, 0);
  S.max_degree = 0;
  {
    int maxdeg[kSegments] = {0};
    for_parts(n_own, kRowGrain, [&](int row0, int row1, int part) {
      int md = 0;
      for (int row = row0; row < row1; ++row) {
        int last = -1, deg = 0;
        for (int s = 0; s < ccnt[row]; ++s) {
          const int col = ccol[S.csell.pos(row, s)];
          if (col < 0 || col == last) continue;
          last = col;
          ++deg;
        }
        rcnt[row] = deg;
        md = std::max(md, deg);
      }
      maxdeg[part] = md;
    });
    for (int t = 0; t < kSegments; ++t) S.max_degree = std::max(S.max_degree, maxdeg[t]);
  }
  S.rsell.build(rcnt);
  const int RS = S.rsell.slots;
  S.r_col.assign(RS, -1);
  S.r_blk.assign(RS, 0);
  std::vector<int>& rptr = S.work.rptr;
  rptr.assign(n_own + 1, 0);
  for (int row = 0; row < n_own; ++row) rptr[row + 1] = rptr[row] + rcnt[row];
  std::vector<int>& nbr = S.work.nbr;  // global columns, CSR
  nbr.resize((size_t)rptr[n_own]);
  for_parts(n_own, kRowGrain, [&](int row0, int row1, int) {
    for (int row = row0; row < row1; ++row) {
      int last = -1, w = rptr[row];
      for (int s = 0; s < ccnt[row]; ++s) {
        const int col = ccol[S.csell.pos(row, s)];
        if (col < 0 || col == last) continue;
        last = col;
        nbr[w++] = col;
      }
    }
  });
  auto ref_cell = [&](int row, int gcol) {  // block cell of reference (row -> gcol)
    const int* b = nbr.data() + rptr[row];
    const int* e = nbr.data() + rptr[row + 1];
    return S.rsell.pos(row, (int)(std::lower_bound(b, e, gcol) - b));
  };
  // The block of pair {a, b} is computed by its writer row: the lo endpoint, or the hi one when lo
  // belongs to another rank.  The writer's own cell holds it in canonical orientation.
  std::vector<int>& rhome = S.work.rhome;  // per reference cell: the cell that holds the canonical block
  rhome.assign(RS, -1);
  S.n_pairs = 0;
  S.nnzo_own = rptr[n_own];
  {
    int pairs[kSegments] = {0};
    for_parts(n_own, kRowGrain, [&](int row0, int row1, int part) {
      int np = 0;
      for (int row = row0; row < row1; ++row) {
        const int g = v0 + row;
        for (int s = 0; s < rcnt[row]; ++s) {
          const int col = nbr[rptr[row] + s];
          const int cell = S.rsell.pos(row, s);
          const int lo = std::min(g, col);
          const bool writer = (g == lo) || !owned(lo);
          S.r_col[cell] = S.local_of(col);
          if (writer) {
            rhome[cell] = cell;
            np++;
          } else {
            rhome[cell] = ref_cell(lo - v0, g);
          }
        }
      }
      pairs[part] = np;
    });
    for (int t = 0; t < kSegments; ++t) S.n_pairs += pairs[t];
  }
  // share or copy, per (slice, slot)
  S.n_block_copies = 0;
  {
    const int ns = (n_own + kSlice - 1) / kSlice;
    int sums[kSegments] = {0};
    for_parts(ns, kSliceGrain, [&](int sl0, int sl1, int part) {
      int away[kSlice], awaycell[kSlice];
      int copies = 0;
      for (int sl = sl0; sl < sl1; ++sl)
        for (int s = 0; s < S.rsell.width[sl]; ++s) {
          int na = 0;
          for (int l = 0; l < kSlice && sl * kSlice + l < n_own; ++l) {
            const int c = S.rsell.base[sl] + s * kSlice + l;
            if (rhome[c] >= 0 && rhome[c] != c) {
              away[na] = rhome[c];
              awaycell[na++] = c;
            }
          }
          if (na && distinct_count(away, na, [](int p) { return p >> 4; }) > kShareMaxLinesBlock) {
            for (int a = 0; a < na; ++a) rhome[awaycell[a]] = awaycell[a];  // this row computes its own copy
            copies += na;
          }
        }
      for (int row = sl0 * kSlice; row < std::min(n_own, sl1 * kSlice); ++row) {
        const int g = v0 + row;
        for (int s = 0; s < rcnt[row]; ++s) {
          const int col = nbr[rptr[row] + s];
          const int cell = S.rsell.pos(row, s);
          // every stored block is canonical (lo orientation): the hi row uses it transposed, whether
          // it reads the lo row's cell or its own copy
          S.r_blk[cell] = (rhome[cell] << 1) | (g != std::min(g, col) ? 1 : 0);
        }
      }
      sums[part] = copies;
    });
    for (int t = 0; t < kSegments; ++t) S.n_block_copies += sums[t];
  }
  // block words of the contributions whose row computes the pair's block into its own cell
  S.c_first.assign(std::max(n_own, 1), 0);
  S.n_merged = 0;
  {
    int merged[kSegments] = {0};
    for_parts(n_own, kRowGrain, [&](int row0, int row1, int part) {
      int nm = 0;
      for (int row = row0; row < row1; ++row) {
        const int g = v0 + row;
        int last = -1, s_ref = -1;
        int first = ccnt[row];
        for (int s = ccnt[row] - 1; s >= 0; --s) {
          const int col = ccol[S.csell.pos(row, s)];
          if (col > g) first = s;
          else break;
        }
        S.c_first[row] = first;
        for (int s = 0; s < ccnt[row]; ++s) {
          const int c = S.csell.pos(row, s);
          const int col = ccol[c];
          if (col < 0) continue;
          const bool head = (col != last);
          if (head) {
            last = col;
            ++s_ref;
          }
          const int lo = std::min(g, col);
          const int cell = S.rsell.pos(row, s_ref);
          if (rhome[cell] != cell) continue;  // the lo row's cell is shared
          S.c_blk[c] = (cell << 2) | ((g != lo) ? 2 : 0) | (head ? 1 : 0);
          if (!head) nm++;
        }
      }
      merged[part] = nm;
    });
    for (int t = 0; t < kSegments; ++t) S.n_merged += merged[t];
  }

  // ---- halo lists.  Rank q needs my row a iff a factor joins a with a row q owns; its ghost list
  //      is then sorted by global index, and so is my send list, so slots line up on both sides.
  S.peers.clear();
  S.send_rows.clear();
  if (world > 1) {
    std::vector<std::vector<int>> send(world);
    for (int q = 0; q < EL; ++q) {
      const int k = perm[loc[q]];
      const bool oi = owned(ei[k]), oj = owned(ej[k]);
      if (oi && !oj) send[S.part.owner(ej[k])].push_back(ei[k] - v0);
      if (oj && !oi) send[S.part.owner(ei[k])].push_back(ej[k] - v0);
    }
    int goff = 0;
    for (int q = 0; q < world; ++q) {
      if (q == rank) continue;
      auto& v = send[q];
      std::sort(v.begin(), v.end());
      v.erase(std::unique(v.begin(), v.end()), v.end());
      const int b = S.part.row_begin(q), e = S.part.row_end(q);
      const int g0 = (int)(std::lower_bound(S.ghosts.begin(), S.ghosts.end(), b) - S.ghosts.begin());
      const int g1 = (int)(std::lower_bound(S.ghosts.begin(), S.ghosts.end(), e) - S.ghosts.begin());
      if (v.empty() && g1 == g0) continue;
      HaloPeer hp;
      hp.rank = q;
      hp.send_off = (int)S.send_rows.size();
      hp.send_cnt = (int)v.size();
      hp.ghost_off = g0;
      hp.ghost_cnt = g1 - g0;
      S.send_rows.insert(S.send_rows.end(), v.begin(), v.end());
      S.peers.push_back(hp);
      goff += hp.ghost_cnt;
    }
    (void)goff;
  }

  // ---- tiles and segments
  S.ntiles = (n_own + kTile - 1) / kTile;
  const int spr = S.part.segs_per_rank();
  for (int s = 0; s <= spr; ++s) {
    const int grow = S.part.seg_row[rank * spr + s];
    S.seg_tile[s] = (grow - v0 + kTile - 1) / kTile;
  }
  S.seg_tile[spr] = S.ntiles;
}

}  // namespace gs
