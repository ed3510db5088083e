// Host-side structure build (K6): the one-off, integer-only work that GTSAM redoes on every
// LevenbergMarquardtOptimizer construction (Ordering / VariableIndex / symbolic factorisation;
// reference src/graph/src/graph.cpp:119).  Output is bit-exact by construction:
//   * canonical edge order  : sort by (min(i,j), max(i,j), direction, insertion index)
//   * contribution lists    : per vertex (block row), every incident factor side, ordered by the
//                             other endpoint (column) then canonical edge order  -> the fixed
//                             summation order of the atomic-free assembly
//   * off-diagonal BSR      : sorted unique columns per row (duplicates / both directions merged)
//   * tiles                 : contiguous row ranges sized for one CTA
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

namespace gs {

constexpr int kAsmChunk = 256;   // contributions staged in shared memory per assembly chunk
constexpr int kAsmRows = 128;    // max rows per assembly tile
constexpr int kSpmvChunk = 512;  // off-diagonal blocks staged per SpMV chunk
constexpr int kSpmvRows = 256;   // max rows per SpMV tile

struct Tile {
  int r0, r1;
};

struct HostStructure {
  int N = 0, E = 0, P = 0;
  int nnzo = 0;  // off-diagonal blocks
  int nc = 0;    // contributions = 2E + P
  std::vector<int> perm;          // canonical edge k -> caller's edge index
  std::vector<int> ci, cj;        // canonical (i, j)
  std::vector<int> crowptr;       // N+1
  std::vector<uint32_t> c_code;   // (index << 2) | kind ; kind 0 = row is id_1, 1 = row is id_2, 2 = prior
  std::vector<int> c_blk;         // (off-diagonal block << 1) | head ; -1 for priors
  std::vector<int> orowptr;       // N+1
  std::vector<int> ocol;          // nnzo
  std::vector<Tile> atiles, stiles;
  int max_row_blocks = 0;

  // sorted unique 3x3 block pattern including the diagonal (the parity hook)
  void full_pattern(std::vector<int32_t>& rowptr, std::vector<int32_t>& colidx) const {
    rowptr.assign(N + 1, 0);
    colidx.resize((size_t)nnzo + N);
    size_t w = 0;
    for (int r = 0; r < N; ++r) {
      rowptr[r] = (int32_t)w;
      int k = orowptr[r];
      const int k1 = orowptr[r + 1];
      while (k < k1 && ocol[k] < r) colidx[w++] = ocol[k++];
      colidx[w++] = r;
      while (k < k1) colidx[w++] = ocol[k++];
    }
    rowptr[N] = (int32_t)w;
  }
};

inline void build_structure(int N, int E, const int* ei, const int* ej, int P, const int* prior_idx, HostStructure& S) {
  S.N = N;
  S.E = E;
  S.P = P;
  S.nc = 2 * E + P;
  // ---- canonical edge order: counting sort by min endpoint, then (max, dir, insertion) inside a group
  std::vector<int> gptr(N + 1, 0);
  for (int k = 0; k < E; ++k) gptr[std::min(ei[k], ej[k]) + 1]++;
  for (int r = 0; r < N; ++r) gptr[r + 1] += gptr[r];
  S.perm.resize(E);
  {
    std::vector<int> fill(gptr.begin(), gptr.end() - 1);
    for (int k = 0; k < E; ++k) S.perm[fill[std::min(ei[k], ej[k])]++] = k;
  }
  for (int r = 0; r < N; ++r) {
    const int a = gptr[r], b = gptr[r + 1];
    if (b - a > 1) {
      std::sort(S.perm.begin() + a, S.perm.begin() + b, [&](int x, int y) {
        const int mx = std::max(ei[x], ej[x]), my = std::max(ei[y], ej[y]);
        if (mx != my) return mx < my;
        const int dx = ei[x] > ej[x], dy = ei[y] > ej[y];
        if (dx != dy) return dx < dy;
        return x < y;
      });
    }
  }
  S.ci.resize(E);
  S.cj.resize(E);
  for (int k = 0; k < E; ++k) {
    S.ci[k] = ei[S.perm[k]];
    S.cj[k] = ej[S.perm[k]];
  }
  // ---- contribution lists per row: priors first, then factor sides in canonical order
  //      (canonical order visits a row's neighbours in ascending column order, see header)
  S.crowptr.assign(N + 1, 0);
  for (int p = 0; p < P; ++p) S.crowptr[prior_idx[p] + 1]++;
  for (int k = 0; k < E; ++k) {
    S.crowptr[S.ci[k] + 1]++;
    S.crowptr[S.cj[k] + 1]++;
  }
  for (int r = 0; r < N; ++r) S.crowptr[r + 1] += S.crowptr[r];
  S.c_code.resize(S.nc);
  std::vector<int> ccol(S.nc);
  {
    std::vector<int> fill(S.crowptr.begin(), S.crowptr.end() - 1);
    for (int p = 0; p < P; ++p) {
      const int c = fill[prior_idx[p]]++;
      S.c_code[c] = ((uint32_t)p << 2) | 2u;
      ccol[c] = -1;
    }
    for (int k = 0; k < E; ++k) {
      const int i = S.ci[k], j = S.cj[k];
      int c = fill[i]++;
      S.c_code[c] = ((uint32_t)k << 2) | 0u;
      ccol[c] = j;
      c = fill[j]++;
      S.c_code[c] = ((uint32_t)k << 2) | 1u;
      ccol[c] = i;
    }
  }
  // ---- off-diagonal BSR pattern: unique columns per row; contribution -> block map
  S.orowptr.assign(N + 1, 0);
  S.c_blk.resize(S.nc);
  S.ocol.clear();
  S.ocol.reserve((size_t)2 * E);
  S.max_row_blocks = 0;
  for (int r = 0; r < N; ++r) {
    S.orowptr[r] = (int)S.ocol.size();
    int last = -1;
    for (int c = S.crowptr[r]; c < S.crowptr[r + 1]; ++c) {
      const int col = ccol[c];
      if (col < 0) {
        S.c_blk[c] = -1;
        continue;
      }
      if (col != last) {
        S.ocol.push_back(col);
        last = col;
        S.c_blk[c] = (int)(((S.ocol.size() - 1) << 1) | 1u);
      } else {
        S.c_blk[c] = (int)((S.ocol.size() - 1) << 1);
      }
    }
    S.max_row_blocks = std::max(S.max_row_blocks, (int)S.ocol.size() - S.orowptr[r]);
  }
  S.nnzo = (int)S.ocol.size();
  S.orowptr[N] = S.nnzo;
  // ---- tiles: greedy contiguous row ranges
  auto make_tiles = [&](const std::vector<int>& ptr, int cap, int rows, std::vector<Tile>& out) {
    out.clear();
    int r0 = 0;
    while (r0 < N) {
      int r1 = r0 + 1;
      while (r1 < N && r1 - r0 < rows && ptr[r1 + 1] - ptr[r0] <= cap) ++r1;
      out.push_back({r0, r1});
      r0 = r1;
    }
  };
  make_tiles(S.crowptr, kAsmChunk, kAsmRows, S.atiles);
  make_tiles(S.orowptr, kSpmvChunk, kSpmvRows, S.stiles);
}

}  // namespace gs
