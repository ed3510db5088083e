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
//   * factor storage         : SoA, in 32-row slices (SELL-32) by "home" row = the smaller endpoint
//                              (the larger one when the smaller is owned by another rank).
//   * contribution lists     : per owned row, every incident factor side ordered by the other
//                              endpoint's global index, then canonical order; SELL-32 laid out so a
//                              warp of 32 consecutive rows reads 32 consecutive words per slot.
//   * H off-diagonal blocks  : ONE 3x3 block per unordered vertex pair, stored in canonical
//                              orientation M = A_lo^T Omega A_hi (lo = smaller global index) at a
//                              SELL-32 position of its home row; the other row uses M^T.
//   * reference lists        : per owned row, its neighbours in global column order as
//                              (local column, block position, transposed?) for the SpMV.
//   * halo lists             : for each peer, the owned rows it needs (send) and the contiguous
//                              range of local ghost slots it fills (recv).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

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

struct LocalStructure {
  Partition part;
  int rank = 0;
  int v0 = 0, v1 = 0;        // owned global rows [v0, v1)
  int n_own = 0, n_ghost = 0;
  int E_glob = 0, E_loc = 0, P_loc = 0;
  int n_pairs = 0;           // stored off-diagonal blocks (unordered pairs with an owned endpoint)
  int nnzo_own = 0;          // off-diagonal blocks of the owned rows of the full (non-symmetric) BSR
  int ntiles = 0;
  int seg_tile[kSegments + 1] = {0};  // local tile ranges of the locally owned segments (first kSegments/P used)
  std::vector<int> ghosts;            // global index of ghost g (sorted)

  // factors, SELL order by home row (padding slots hold a harmless dummy)
  Sell fsell;
  std::vector<int> f_src;             // storage slot -> caller's factor index, -1 padding
  std::vector<int> f_i, f_j;          // local vertex index of id_1 / id_2 per slot (-1 padding)
  std::vector<char> f_count;          // 1: this rank counts the factor's chi2 (its lo endpoint is owned)
  // priors on owned rows
  std::vector<int> pr_src;            // local prior -> caller's prior index
  std::vector<int> pr_row;            // local row
  // contribution lists
  Sell csell;
  std::vector<uint32_t> c_code;
  std::vector<int> c_blk;
  std::vector<int> c_col;             // local index of the other vertex of each contribution (0 for priors/padding)
  // H blocks
  Sell bsell;
  // reference lists
  Sell rsell;
  std::vector<int> r_col, r_blk;
  // halo
  std::vector<HaloPeer> peers;
  std::vector<int> send_rows;         // local owned row indices, grouped by peer, each group sorted
  int max_degree = 0;

  int local_of(int g) const {
    if (g >= v0 && g < v1) return g - v0;
    const auto it = std::lower_bound(ghosts.begin(), ghosts.end(), g);
    return n_own + (int)(it - ghosts.begin());
  }
};

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
  std::vector<int> loc;  // canonical positions (index into perm) of local factors
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

  // ---- factor storage: SELL by home row
  std::vector<int> home(EL);
  std::vector<int> fcnt(n_own, 0);
  for (int q = 0; q < EL; ++q) {
    const int k = perm[loc[q]];
    const int lo = std::min(ei[k], ej[k]), hi = std::max(ei[k], ej[k]);
    home[q] = (owned(lo) ? lo : hi) - v0;
    fcnt[home[q]]++;
  }
  S.fsell.build(fcnt);
  S.f_src.assign(S.fsell.slots, -1);
  S.f_i.assign(S.fsell.slots, -1);
  S.f_j.assign(S.fsell.slots, -1);
  S.f_count.assign(S.fsell.slots, 0);
  std::vector<int> fpos(EL);
  {
    std::vector<int> fill(n_own, 0);
    for (int q = 0; q < EL; ++q) {
      const int k = perm[loc[q]];
      const int p = S.fsell.pos(home[q], fill[home[q]]++);
      fpos[q] = p;
      S.f_src[p] = k;
      S.f_i[p] = S.local_of(ei[k]);
      S.f_j[p] = S.local_of(ej[k]);
      S.f_count[p] = owned(std::min(ei[k], ej[k])) ? 1 : 0;
    }
  }

  // ---- priors on owned rows (caller's order)
  S.pr_src.clear();
  S.pr_row.clear();
  for (int p = 0; p < P; ++p)
    if (owned(prior_idx[p])) S.pr_src.push_back(p);
  // ascending row, caller's order within a row (the chi2 pass looks a row's priors up by bisection)
  std::stable_sort(S.pr_src.begin(), S.pr_src.end(), [&](int a, int b) { return prior_idx[a] < prior_idx[b]; });
  for (int p : S.pr_src) S.pr_row.push_back(prior_idx[p] - v0);
  S.P_loc = (int)S.pr_src.size();

  // ---- contribution lists: priors first, then factor sides in canonical order (which visits a
  //      row's neighbours in ascending global column order)
  std::vector<int> ccnt(n_own, 0);
  for (int p = 0; p < S.P_loc; ++p) ccnt[S.pr_row[p]]++;
  for (int q = 0; q < EL; ++q) {
    const int k = perm[loc[q]];
    if (owned(ei[k])) ccnt[ei[k] - v0]++;
    if (owned(ej[k])) ccnt[ej[k] - v0]++;
  }
  S.csell.build(ccnt);
  S.c_code.assign(S.csell.slots, kPadCode);
  S.c_blk.assign(S.csell.slots, -1);
  std::vector<int> ccol(S.csell.slots, -1);  // global column of each contribution (-1 prior)
  {
    std::vector<int> fill(n_own, 0);
    for (int p = 0; p < S.P_loc; ++p) {
      const int row = S.pr_row[p];
      const int c = S.csell.pos(row, fill[row]++);
      S.c_code[c] = ((uint32_t)p << 3) | 4u | 2u;
      ccol[c] = -1;
    }
    for (int q = 0; q < EL; ++q) {
      const int k = perm[loc[q]];
      const int gi = ei[k], gj = ej[k];
      const int lo = std::min(gi, gj);
      if (owned(gi)) {
        const int row = gi - v0;
        const int c = S.csell.pos(row, fill[row]++);
        S.c_code[c] = ((uint32_t)fpos[q] << 3) | (gi == lo ? 4u : 0u) | 0u;
        ccol[c] = gj;
      }
      if (owned(gj)) {
        const int row = gj - v0;
        const int c = S.csell.pos(row, fill[row]++);
        S.c_code[c] = ((uint32_t)fpos[q] << 3) | (gj == lo ? 4u : 0u) | 1u;
        ccol[c] = gi;
      }
    }
  }

  S.c_col.assign(S.csell.slots, 0);
  for (int c = 0; c < S.csell.slots; ++c)
    if (ccol[c] >= 0) S.c_col[c] = S.local_of(ccol[c]);

  // ---- H blocks: one per unordered pair, homed like the factors; block words of the writer rows;
  //      reference triples (global col, local col, block word) per owned row
  std::vector<int> bcnt(n_own, 0), rcnt(n_own, 0);
  S.max_degree = 0;
  for (int row = 0; row < n_own; ++row) {
    const int g = v0 + row;
    int last = -1, deg = 0;
    for (int s = 0; s < ccnt[row]; ++s) {
      const int col = ccol[S.csell.pos(row, s)];
      if (col < 0 || col == last) continue;
      last = col;
      ++deg;
      const int lo = std::min(g, col);
      const bool writer = (g == lo) || !owned(lo);
      if (writer) bcnt[row]++;
    }
    rcnt[row] = deg;
    S.max_degree = std::max(S.max_degree, deg);
  }
  S.bsell.build(bcnt);
  S.rsell.build(rcnt);
  S.n_pairs = 0;
  S.nnzo_own = 0;
  for (int row = 0; row < n_own; ++row) {
    S.n_pairs += bcnt[row];
    S.nnzo_own += rcnt[row];
  }
  S.r_col.assign(S.rsell.slots, -1);
  S.r_blk.assign(S.rsell.slots, 0);
  struct Ref {
    int gcol, lcol, word;
  };
  std::vector<int> rptr(n_own + 1, 0);
  for (int row = 0; row < n_own; ++row) rptr[row + 1] = rptr[row] + rcnt[row];
  std::vector<Ref> refs((size_t)rptr[n_own]);
  {
    std::vector<int> rfill(rptr.begin(), rptr.end() - 1);
    for (int row = 0; row < n_own; ++row) {
      const int g = v0 + row;
      int last = -1, nb = 0, bpos = -1;
      for (int s = 0; s < ccnt[row]; ++s) {
        const int c = S.csell.pos(row, s);
        const int col = ccol[c];
        if (col < 0) continue;
        const int lo = std::min(g, col);
        const bool writer = (g == lo) || !owned(lo);
        if (!writer) continue;
        const bool head = (col != last);
        if (head) {
          last = col;
          bpos = S.bsell.pos(row, nb++);
          const int lcol = S.local_of(col);
          const int hi_row = (g != lo);  // this row is the hi endpoint -> it uses M^T
          refs[rfill[row]++] = {col, lcol, (bpos << 1) | hi_row};
          if (owned(col)) {  // the other endpoint is an owned row too: it references the same block
            const int orow = col - v0;
            refs[rfill[orow]++] = {g, row, (bpos << 1) | (hi_row ? 0 : 1)};
          }
        }
        S.c_blk[c] = (bpos << 2) | ((g != lo) ? 2 : 0) | (head ? 1 : 0);
      }
    }
  }
  for (int row = 0; row < n_own; ++row) {
    std::sort(refs.begin() + rptr[row], refs.begin() + rptr[row + 1], [](const Ref& a, const Ref& b) { return a.gcol < b.gcol; });
    for (int s = 0; s < rcnt[row]; ++s) {
      const Ref& f = refs[rptr[row] + s];
      const int p = S.rsell.pos(row, s);
      S.r_col[p] = f.lcol;
      S.r_blk[p] = f.word;
    }
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
