// Per-row bodies of the Gauss-Newton / Levenberg-Marquardt kernels.  One thread owns one block row
// (one vertex): it walks the row's SELL-32 list, so the 32 lanes of a warp read 32 consecutive
// words per slot (coalesced), every output has exactly one writer (no atomics) and every sum has
// one fixed order.  The functions are __host__ __device__: the kernels in kernels.cuh call them per
// thread, and the test-only emulator (emu.cu) calls the very same code per row on the CPU.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>
#include "se2.cuh"
#include "structure.hpp"

namespace gs {

template <typename T>
GS_HD T ldg(const T* p) {
#ifdef __CUDA_ARCH__
  return __ldg(p);
#else
  return *p;
#endif
}

// a pose is one aligned 32-byte sector: two 16-byte read-only loads
GS_HD double4 ldg(const double4* p) {
#ifdef __CUDA_ARCH__
  const double2 a = __ldg(reinterpret_cast<const double2*>(p));
  const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
  return make_double4(a.x, a.y, b.x, b.y);
#else
  return *p;
#endif
}

struct SellView {
  const int* base;   // per slice
  const int* width;  // per slice
};

// ---- storage layouts of the two big per-cell arrays (compile-time; results do not depend on them) -------------
// GS_H_LAYOUT  0: nine planes of hstride doubles (entry k of block cell q at k*hstride + q)
//              1: slice-major: the nine entries of the 32 cells of one (slice, slot) are adjacent, 9 x 256 B
//              2: slice-major in pairs: entries (0,1) (2,3) (4,5) (6,7) of a cell are adjacent (16-byte accesses,
//                 4 x 512 B per group), entry 8 follows (256 B): five accesses per block instead of nine
// GS_F_LAYOUT  0: five planes of fstride double2 (component c of factor cell q at c*fstride + q)
//              1: slice-major: the five components of the 32 cells of one (slice, slot) are adjacent, 5 x 512 B
// Cells come in groups of 32 (SELL-32 lists), so both strides are multiples of 32.
#ifndef GS_H_LAYOUT
#define GS_H_LAYOUT 1   // measured on C3 / C4 (B200): k_spmv 100.5 -> 94.5 us / 144.8 -> 134.5 us against layout 0
#endif
#ifndef GS_F_LAYOUT
#define GS_F_LAYOUT 1   // k_assemble 179.8 -> 177 us on C3, 319.8 -> 312.9 us on C4
#endif
#ifndef GS_SPMV_PRELOAD
#define GS_SPMV_PRELOAD 1   // 1: the row's diagonal block is requested before the slot loop (C4: 143 -> 135 us)
#endif

// offset of entry 0 of block cell q, and the distance between consecutive entries of one block
GS_HD size_t h_base(int q) {
#if GS_H_LAYOUT >= 1
  return (size_t)(q >> 5) * (9 * kSlice) + (size_t)(q & 31);
#else
  return (size_t)q;
#endif
}
GS_HD size_t h_step(size_t hstride) {
#if GS_H_LAYOUT >= 1
  (void)hstride;
  return kSlice;
#else
  return hstride;
#endif
}
// offset of component 0 of factor cell q, and the distance between consecutive components
GS_HD size_t f_base(int q) {
#if GS_F_LAYOUT == 1
  return (size_t)(q >> 5) * (5 * kSlice) + (size_t)(q & 31);
#else
  return (size_t)q;
#endif
}
GS_HD size_t f_step(size_t fstride) {
#if GS_F_LAYOUT == 1
  (void)fstride;
  return kSlice;
#else
  return fstride;
#endif
}

constexpr int kNoCountBit = 1 << 30;   // set in ij.x of a stored factor whose chi2 another rank counts
constexpr int kIdxMask = kNoCountBit - 1;

struct FactorView {
  const int2* ij;       // per factor cell: local indices of id_1 (| kNoCountBit), id_2 ; (-1,-1) = no data
  const double2* data;  // five double2 components per cell, at f_base(cell) + c * f_step(fstride):
                        //   0 measured translation, 1 (cos, sin) of the measured rotation,
                        //   2..4 information, upper triangle 00 01 | 02 11 | 12 22
  size_t fstride;
};

struct FactorData {
  double2 zt, zr, o01, o23, o45;
};

GS_HD FactorData load_factor(const FactorView& f, int cell) {
  const double2* p = f.data + f_base(cell);
  const size_t st = f_step(f.fstride);
  FactorData d;
  d.zt = ldg(p);
  d.zr = ldg(p + st);
  d.o01 = ldg(p + 2 * st);
  d.o23 = ldg(p + 3 * st);
  d.o45 = ldg(p + 4 * st);
  return d;
}

// host staging of the factor cells in the device layout (the structure build's f_src / f_i / f_j / f_count say
// which cells hold data); `cells` = csell.slots, a multiple of 32
// one cell of the staged factor arrays (ij: fstride entries, data: 5 * fstride), raw storage
inline void factor_cell_default(int2* ij, double2* data, size_t fstride, int q) {
  ij[q] = make_int2(-1, -1);
  double2* p = data + f_base(q);
  const size_t st = f_step(fstride);
  p[0] = make_double2(0.0, 0.0);
  p[st] = make_double2(1.0, 0.0);   // unit rotation
  p[2 * st] = make_double2(0.0, 0.0);
  p[3 * st] = make_double2(0.0, 0.0);
  p[4 * st] = make_double2(0.0, 0.0);
}

// z = (dx, dy, dtheta), om = information upper triangle (6)
inline void factor_cell_set(int2* ij, double2* data, size_t fstride, int q, int i, int j, const double* z, const double* om) {
  ij[q] = make_int2(i, j);
  double2* p = data + f_base(q);
  const size_t st = f_step(fstride);
  p[0] = make_double2(z[0], z[1]);
  p[st] = make_double2(std::cos(z[2]), std::sin(z[2]));
  p[2 * st] = make_double2(om[0], om[1]);
  p[3 * st] = make_double2(om[2], om[3]);
  p[4 * st] = make_double2(om[4], om[5]);
}

inline size_t factor_stride(int cells) { return (size_t)((std::max(cells, 1) + kSlice - 1) / kSlice) * kSlice; }

struct FactorStage {
  std::vector<int2> ij;
  std::vector<double2> data;
  size_t fstride = kSlice;
  void init(int cells) {
    fstride = factor_stride(cells);
    ij.resize(fstride);
    data.resize(5 * fstride);
    for (size_t q = 0; q < fstride; ++q) factor_cell_default(ij.data(), data.data(), fstride, (int)q);
  }
  void set(int q, int i, int j, const double* z, const double* om) {
    factor_cell_set(ij.data(), data.data(), fstride, q, i, j, z, om);
  }
};

struct PriorView {
  int n;                 // priors on this rank's rows
  const int* row;        // local row, ascending
  const double4* pose;   // (x, y, cos, sin)
  const double* om;      // 6 per prior
};

// ------------------------------------------------------------------------------------------------
// K1+K2: linearise every factor side of row r and reduce it into the row's diagonal block, its
// gradient, and the off-diagonal blocks this row is the writer of.
// ------------------------------------------------------------------------------------------------
struct AsmView {
  SellView cs;
  const int4* c;         // per contribution: (code, block word, local index of the other vertex, unused);
                         // code == kPadCode: padding slot
  const int* c_first;    // per row: first slot of its counting cells (factors with this row as lo endpoint)
  FactorView f;
  PriorView pr;
  const double4* pose;   // n_own + n_ghost
  double* Ho;            // 9 planes of hstride
  size_t hstride;
  double* Hd;            // 6 planes of n_own
  double* b;             // 3 * n_own
  int n_own;
};

// canonical off-diagonal block M = A_lo^T Omega A_hi, row-major.  One call site, so the two ranks
// that both need the block of a cut factor produce the same bits.
GS_HD void off_block(const Sym3& O, const Mat3& Alo, const Mat3& Ahi, double M[9]) {
  double W[9];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    W[0 + c] = O.a00 * Alo.m[0 + c] + O.a01 * Alo.m[3 + c] + O.a02 * Alo.m[6 + c];
    W[3 + c] = O.a01 * Alo.m[0 + c] + O.a11 * Alo.m[3 + c] + O.a12 * Alo.m[6 + c];
    W[6 + c] = O.a02 * Alo.m[0 + c] + O.a12 * Alo.m[3 + c] + O.a22 * Alo.m[6 + c];
  }
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) M[3 * r + c] = W[0 + r] * Ahi.m[0 + c] + W[3 + r] * Ahi.m[3 + c] + W[6 + r] * Ahi.m[6 + c];
}

// diagonal and gradient pieces of one factor side: diag = Ar^T Omega Ar (sym 6), bv = Ar^T Omega e
GS_HD void diag_pieces(const Sym3& O, const Mat3& Ar, const double e[3], double diag[6], double bv[3]) {
  double W[9];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    W[0 + c] = O.a00 * Ar.m[0 + c] + O.a01 * Ar.m[3 + c] + O.a02 * Ar.m[6 + c];
    W[3 + c] = O.a01 * Ar.m[0 + c] + O.a11 * Ar.m[3 + c] + O.a12 * Ar.m[6 + c];
    W[6 + c] = O.a02 * Ar.m[0 + c] + O.a12 * Ar.m[3 + c] + O.a22 * Ar.m[6 + c];
  }
  diag[0] = Ar.m[0] * W[0] + Ar.m[3] * W[3] + Ar.m[6] * W[6];
  diag[1] = Ar.m[0] * W[1] + Ar.m[3] * W[4] + Ar.m[6] * W[7];
  diag[2] = Ar.m[0] * W[2] + Ar.m[3] * W[5] + Ar.m[6] * W[8];
  diag[3] = Ar.m[1] * W[1] + Ar.m[4] * W[4] + Ar.m[7] * W[7];
  diag[4] = Ar.m[1] * W[2] + Ar.m[4] * W[5] + Ar.m[7] * W[8];
  diag[5] = Ar.m[2] * W[2] + Ar.m[5] * W[5] + Ar.m[8] * W[8];
#pragma unroll
  for (int r = 0; r < 3; ++r) bv[r] = W[0 + r] * e[0] + W[3 + r] * e[1] + W[6 + r] * e[2];
}

GS_HD void store_block(double* Ho, size_t hstride, int pos, const double M[9]) {
#if GS_H_LAYOUT == 2
  (void)hstride;
  double* g = Ho + (size_t)(pos >> 5) * (9 * kSlice);
  const int l = pos & 31;
  double2* g2 = reinterpret_cast<double2*>(g);
#pragma unroll
  for (int m = 0; m < 4; ++m) g2[m * kSlice + l] = make_double2(M[2 * m], M[2 * m + 1]);
  g[8 * kSlice + l] = M[8];
#else
  double* hp = Ho + h_base(pos);
  const size_t st = h_step(hstride);
#pragma unroll
  for (int k = 0; k < 9; ++k) hp[(size_t)k * st] = M[k];
#endif
}

GS_HD void load_block(const double* Ho, size_t hstride, int pos, double h[9]) {
#if GS_H_LAYOUT == 2
  (void)hstride;
  const double* g = Ho + (size_t)(pos >> 5) * (9 * kSlice);
  const int l = pos & 31;
  const double2* g2 = reinterpret_cast<const double2*>(g);
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const double2 v = ldg(g2 + m * kSlice + l);
    h[2 * m] = v.x;
    h[2 * m + 1] = v.y;
  }
  h[8] = ldg(g + 8 * kSlice + l);
#else
  const double* hp = Ho + h_base(pos);
  const size_t st = h_step(hstride);
#pragma unroll
  for (int k = 0; k < 9; ++k) h[k] = ldg(hp + (size_t)k * st);
#endif
}

// The same block as the row that uses it needs it: transposed when `tr` (the hi row of the pair reads the lo row's
// canonical block).  The transposition is taken in the ADDRESSES (entries 1<->3, 2<->6, 5<->7 swap), not by selecting
// between loaded values: the nine loads are independent of each other and of the arithmetic, so they all go out
// before the first use instead of in four dependent groups (the selects forced register reuse between groups).
GS_HD void load_block_as(const double* Ho, size_t hstride, int pos, bool tr, double a[9]) {
#if GS_H_LAYOUT == 2
  double h[9];
  load_block(Ho, hstride, pos, h);
  a[0] = h[0]; a[4] = h[4]; a[8] = h[8];
  a[1] = tr ? h[3] : h[1]; a[3] = tr ? h[1] : h[3];
  a[2] = tr ? h[6] : h[2]; a[6] = tr ? h[2] : h[6];
  a[5] = tr ? h[7] : h[5]; a[7] = tr ? h[5] : h[7];
#else
  const double* hp = Ho + h_base(pos);
  const size_t st = h_step(hstride);
  const size_t s1 = tr ? 3 * st : st, s3 = tr ? st : 3 * st;
  const size_t s2 = tr ? 6 * st : 2 * st, s6 = tr ? 2 * st : 6 * st;
  const size_t s5 = tr ? 7 * st : 5 * st, s7 = tr ? 5 * st : 7 * st;
  a[0] = ldg(hp);
  a[1] = ldg(hp + s1);
  a[2] = ldg(hp + s2);
  a[3] = ldg(hp + s3);
  a[4] = ldg(hp + 4 * st);
  a[5] = ldg(hp + s5);
  a[6] = ldg(hp + s6);
  a[7] = ldg(hp + s7);
  a[8] = ldg(hp + 8 * st);
#endif
}

// entry (r, c) of the stored block cell `pos` of a host copy of Ho, optionally transposed
inline double h_entry(const double* Ho, size_t hstride, int pos, int r, int c, bool tr) {
  double h[9];
  load_block(Ho, hstride, pos, h);
  return h[tr ? 3 * c + r : 3 * r + c];
}

// One factor side under GTSAM's default Jacobians (A_i = [[-c,-s,dt1],[s,-c,dt2],[0,0,-1]], A_j = I):
// everything the row needs follows from W = Omega A_i —
//   row is id_1:  diag += A_i^T W,  b += W^T e        row is id_2:  diag += Omega,  b += Omega e
//   canonical off-diagonal block  M = A_lo^T Omega A_hi  =  W^T if lo is id_1, W if lo is id_2
// W is formed at ONE place, so the two ranks that both store the block of a cut factor get the same bits.
struct SideDefault {
  double W[9];
  double e[3];
  double c, s, dt1, dt2;
};

GS_HD void side_default(const double4 pi, const double4 pj, const double2 zt, const double2 zr, const Sym3& O,
                        SideDefault& S) {
  const double ci = pi.z, si = pi.w, cj = pj.z, sj = pj.w;
  const double dx = pj.x - pi.x, dy = pj.y - pi.y;
  const double hx = ci * dx + si * dy;
  const double hy = -si * dx + ci * dy;
  double c = ci * cj + si * sj;
  double s = -si * cj + ci * sj;
  rot2_normalize(c, s);  // Rot2::atan2(s, c)
  const double cm = zr.x, sm = zr.y;
  const double rx = hx - zt.x, ry = hy - zt.y;
  S.e[0] = cm * rx + sm * ry;
  S.e[1] = -sm * rx + cm * ry;
  S.e[2] = atan2(s * cm - c * sm, c * cm + s * sm);
  const double dt1 = -sj * dx + cj * dy;
  const double dt2 = -cj * dx - sj * dy;
  S.c = c;
  S.s = s;
  S.dt1 = dt1;
  S.dt2 = dt2;
  const double r0[3] = {O.a00, O.a01, O.a02}, r1[3] = {O.a01, O.a11, O.a12}, r2[3] = {O.a02, O.a12, O.a22};
  const double* rows[3] = {r0, r1, r2};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double a = rows[k][0], b = rows[k][1], d = rows[k][2];
    S.W[3 * k + 0] = s * b - c * a;
    S.W[3 * k + 1] = -(s * a) - c * b;
    S.W[3 * k + 2] = (dt1 * a + dt2 * b) - d;
  }
}

// running state of one row during assembly
struct AsmAcc {
  double dg[6];   // diagonal block, symmetric
  double bv[3];   // gradient
  double acc[9];  // off-diagonal block being summed (a run of factors of one vertex pair)
  double chi2;
  int open;       // cell of the block in acc, -1 none
};

GS_HD void asm_begin(AsmAcc& A) {
#pragma unroll
  for (int k = 0; k < 6; ++k) A.dg[k] = 0.0;
#pragma unroll
  for (int k = 0; k < 3; ++k) A.bv[k] = 0.0;
  A.chi2 = 0.0;
  A.open = -1;
}

// one BetweenFactor side: code/cb are the contribution's words, prow/pcol the poses of the row's and the
// other vertex, (zt, zr, o01, o23, o45) the factor
// NODUP: the structure has no two factors between the same pair of vertices (LocalStructure::n_merged == 0), so
// every run is one factor long and its block is stored as soon as it is formed: no run accumulator (18 registers)
template <int MODE, bool NODUP = false>
GS_HD void asm_between(AsmAcc& A, uint32_t code, int cb, const double4 prow, const double4 pcol, const double2 zt,
                       const double2 zr, const double2 o01, const double2 o23, const double2 o45, double* Ho,
                       size_t hstride) {
  const Sym3 O{o01.x, o01.y, o23.x, o23.y, o45.x, o45.y};
  const bool row_is_i = ((code & 3u) == 0);
  const double4 pi = row_is_i ? prow : pcol, pj = row_is_i ? pcol : prow;
  // row is the hi endpoint of the pair (cb & 2): the lo Jacobian is the other vertex's
  const bool lo_is_i = row_is_i != ((cb & 2) != 0);
  double M[9];
  if (MODE == 0) {
    SideDefault S;
    side_default(pi, pj, zt, zr, O, S);
    const double* W = S.W;
    if (row_is_i) {
      A.dg[0] += S.s * W[3] - S.c * W[0];
      A.dg[1] += S.s * W[4] - S.c * W[1];
      A.dg[2] += S.s * W[5] - S.c * W[2];
      A.dg[3] += -(S.s * W[1]) - S.c * W[4];
      A.dg[4] += -(S.s * W[2]) - S.c * W[5];
      A.dg[5] += (S.dt1 * W[2] + S.dt2 * W[5]) - W[8];
#pragma unroll
      for (int k = 0; k < 3; ++k) A.bv[k] += (W[k] * S.e[0] + W[3 + k] * S.e[1]) + W[6 + k] * S.e[2];
    } else {
      A.dg[0] += O.a00; A.dg[1] += O.a01; A.dg[2] += O.a02;
      A.dg[3] += O.a11; A.dg[4] += O.a12; A.dg[5] += O.a22;
      A.bv[0] += (O.a00 * S.e[0] + O.a01 * S.e[1]) + O.a02 * S.e[2];
      A.bv[1] += (O.a01 * S.e[0] + O.a11 * S.e[1]) + O.a12 * S.e[2];
      A.bv[2] += (O.a02 * S.e[0] + O.a12 * S.e[1]) + O.a22 * S.e[2];
    }
    if (code & 4u) A.chi2 += quad_form(O, S.e);
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
      for (int t = 0; t < 3; ++t) M[3 * q + t] = lo_is_i ? W[3 * t + q] : W[3 * q + t];
  } else {
    BetweenLin L;
    between_linearize<MODE>(pi, pj, zt, zr, L);
    double d6[6], b3[3];
    diag_pieces(O, row_is_i ? L.Ai : L.Aj, L.e, d6, b3);
#pragma unroll
    for (int k = 0; k < 6; ++k) A.dg[k] += d6[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) A.bv[k] += b3[k];
    if (code & 4u) A.chi2 += quad_form(O, L.e);
    if (cb >= 0) off_block(O, lo_is_i ? L.Ai : L.Aj, lo_is_i ? L.Aj : L.Ai, M);
  }
  if (NODUP) {
    if (cb >= 0) store_block(Ho, hstride, cb >> 2, M);
  } else if (cb >= 0) {
    if (cb & 1) {
      if (A.open >= 0) store_block(Ho, hstride, A.open, A.acc);
      A.open = cb >> 2;
#pragma unroll
      for (int k = 0; k < 9; ++k) A.acc[k] = M[k];
    } else {
#pragma unroll
      for (int k = 0; k < 9; ++k) A.acc[k] += M[k];
    }
  }
}

template <int MODE>
GS_HD void asm_prior(AsmAcc& A, const double4 prow, const double4 pp, const double* om) {
  const Sym3 O{om[0], om[1], om[2], om[3], om[4], om[5]};
  double e[3], d6[6], b3[3];
  Mat3 J;
  prior_linearize<MODE>(prow, pp, e, J);
  diag_pieces(O, J, e, d6, b3);
  A.chi2 += quad_form(O, e);
#pragma unroll
  for (int k = 0; k < 6; ++k) A.dg[k] += d6[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) A.bv[k] += b3[k];
}

template <bool NODUP = false>
GS_HD double asm_end(AsmAcc& A, const AsmView& a, int r) {
  if (!NODUP && A.open >= 0) store_block(a.Ho, a.hstride, A.open, A.acc);
#pragma unroll
  for (int k = 0; k < 6; ++k) a.Hd[(size_t)k * a.n_own + r] = A.dg[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) a.b[3 * (size_t)r + k] = A.bv[k];
  return A.chi2;
}

// returns the row's chi2 share (factors it is the counting side of, and its priors)
template <int MODE, bool NODUP = false>
GS_HD double assemble_row(const AsmView& a, int r) {
  const int sl = r / kSlice, lane = r % kSlice;
  const int base = a.cs.base[sl] + lane, w = a.cs.width[sl];
  AsmAcc A;
  asm_begin(A);
  const double4 prow = ldg(a.pose + r);
  // w is the same for the 32 rows of a slice.  The next slot's words are fetched while this slot's
  // factor and pose are in flight; each slot is then one dependent round trip (factor + other pose).
  int4 cw = w > 0 ? ldg(a.c + base) : make_int4((int)kPadCode, -1, 0, -1);
  for (int s = 0; s < w; ++s) {
    const int4 cur = cw;
    if (s + 1 < w) cw = ldg(a.c + base + (s + 1) * kSlice);
    const uint32_t code = (uint32_t)cur.x;
    if (code == kPadCode) continue;
    const int idx = (int)(code >> 3);
    if ((code & 3u) < 2) {
      const FactorData fd = load_factor(a.f, idx);
      const double4 pcol = ldg(a.pose + cur.z);
      asm_between<MODE, NODUP>(A, code, cur.y, prow, pcol, fd.zt, fd.zr, fd.o01, fd.o23, fd.o45, a.Ho, a.hstride);
    } else {
      asm_prior<MODE>(A, prow, ldg(a.pr.pose + idx), a.pr.om + 6 * idx);
    }
  }
  return asm_end<NODUP>(A, a, r);
}

// one counting cell of k_chi2 in flight: the factor and the poses of its two vertices
struct Chi2Stage {
  FactorData fd;
  double4 pi, pj;
  bool ok;
};

GS_HD void chi2_request(const AsmView& a, int cell, const int2 ij, Chi2Stage& S) {
  S.ok = ij.x >= 0 && !(ij.x & kNoCountBit);
  if (S.ok) {
    S.fd = load_factor(a.f, cell);
    S.pi = ldg(a.pose + ij.x);
    S.pj = ldg(a.pose + ij.y);
  }
}

GS_HD double chi2_value(const Chi2Stage& S) {
  double e[3];
  between_error(S.pi, S.pj, S.fd.zt, S.fd.zr, e);
  const Sym3 O{S.fd.o01.x, S.fd.o01.y, S.fd.o23.x, S.fd.o23.y, S.fd.o45.x, S.fd.o45.y};
  return quad_form(O, e);
}

// graph.error contribution of row r at `pose`: its priors, then the factors whose counting side is
// one of its own cells (same set and order as assemble_row's chi2).  The factor arrays are read
// directly in cell order: 32 consecutive rows read 32 consecutive cells.
// PIPE: how far the slot loop runs ahead of its arithmetic.  A slot is two dependent round trips (the cell's vertex
// pair, then the two poses and the factor) followed by ~150 dependent FP64 instructions (atan2).
//   0  one slot after the other (what the plain loop compiles to): 2 n round trips per row of n factors
//   1  the pair of slot k+1 is requested before slot k's data: n + 1 round trips, two more registers
//   2  the pair of slot k+2 AND the data of slot k+1 are requested before the arithmetic of slot k: the thread always
//      has a factor in flight, at 104 registers (two CTAs per SM instead of four)
// Same factors in the same order: same bits.  Measured (profiles/r2_summary.md): rows of three factors (C3) are
// fastest at 1, rows of twenty (C4) at 2; the library picks by the mean number of factors per row.
template <int PIPE>
GS_HD double chi2_row(const AsmView& a, int r) {
  double chi2 = 0.0;
  {  // priors of this row: lower_bound in the ascending row list
    int lo = 0, hi = a.pr.n;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (a.pr.row[mid] < r) lo = mid + 1;
      else hi = mid;
    }
    for (int k = lo; k < a.pr.n && a.pr.row[k] == r; ++k) {
      const double* om = a.pr.om + 6 * k;
      const Sym3 O{om[0], om[1], om[2], om[3], om[4], om[5]};
      double e[3];
      Mat3 A;
      prior_linearize<0>(ldg(a.pose + r), ldg(a.pr.pose + k), e, A);
      chi2 += quad_form(O, e);
    }
  }
  const int sl = r / kSlice, lane = r % kSlice;
  const int base = a.cs.base[sl] + lane, w = a.cs.width[sl];
  const int first = a.c_first[r];
  const int n = w - first;            // slots first .. w-1 are the row's counting cells
  if (n <= 0) return chi2;
  const int2 none = make_int2(-1, -1);
  const int cell0 = base + first * kSlice;
  const int2* cell_ij = a.f.ij + cell0;
  if (PIPE == 2) {
    int2 ij_next = n > 1 ? ldg(cell_ij + kSlice) : none;
    Chi2Stage cur;
    chi2_request(a, cell0, ldg(cell_ij), cur);
#pragma unroll 1
    for (int k = 0; k < n; ++k) {
      const int2 ij_after = k + 2 < n ? ldg(cell_ij + (k + 2) * kSlice) : none;
      Chi2Stage nxt;
      chi2_request(a, cell0 + (k + 1) * kSlice, ij_next, nxt);   // ij_next is (-1, -1) past the end: nothing requested
      if (cur.ok) chi2 += chi2_value(cur);
      cur = nxt;
      ij_next = ij_after;
    }
  } else if (PIPE == 1) {
    int2 ij = ldg(cell_ij);
#pragma unroll 1
    for (int k = 0; k < n; ++k) {
      const int2 ij_next = k + 1 < n ? ldg(cell_ij + (k + 1) * kSlice) : none;
      Chi2Stage cur;
      chi2_request(a, cell0 + k * kSlice, ij, cur);
      if (cur.ok) chi2 += chi2_value(cur);
      ij = ij_next;
    }
  } else {
#pragma unroll 4
    for (int k = 0; k < n; ++k) {
      Chi2Stage cur;
      chi2_request(a, cell0 + k * kSlice, ldg(cell_ij + k * kSlice), cur);
      if (cur.ok) chi2 += chi2_value(cur);
    }
  }
  return chi2;
}

// ------------------------------------------------------------------------------------------------
// K3: y_r = sum_c H_rc p_c (global column order) + H_rr p_r + damp p_r ; returns p_r . y_r
// ------------------------------------------------------------------------------------------------
struct SpmvView {
  SellView rs;
  const int2* refs;      // (local column, block word) per neighbour; column < 0: padding slot
  const double* Ho;
  size_t hstride;
  const double* Hd;      // 6 planes of n_own
  const double* dscale;  // nullptr: lambda*I ; else lambda*diag(dscale), 3 per row
  int n_own;
};

// one neighbour block: y += A p_c, A = the block as this row uses it (load_block_as), row-major
GS_HD void spmv_slot(const double a[9], double px, double py, double pz, double& y0, double& y1, double& y2) {
  y0 += (a[0] * px + a[1] * py) + a[2] * pz;
  y1 += (a[3] * px + a[4] * py) + a[5] * pz;
  y2 += (a[6] * px + a[7] * py) + a[8] * pz;
}

// diagonal block, damping, store, and the row's share of p.y ; d = (00 01 02 11 12 22), l = damping per component
GS_HD double spmv_finish(const double d[6], const double l[3], double px, double py, double pz, double y0, double y1,
                         double y2, double* y, size_t rr) {
  y0 = (y0 + ((d[0] * px + d[1] * py) + d[2] * pz)) + l[0] * px;
  y1 = (y1 + ((d[1] * px + d[3] * py) + d[4] * pz)) + l[1] * py;
  y2 = (y2 + ((d[2] * px + d[4] * py) + d[5] * pz)) + l[2] * pz;
  y[3 * rr] = y0;
  y[3 * rr + 1] = y1;
  y[3 * rr + 2] = y2;
  return (px * y0 + py * y1) + pz * y2;
}

// the PCG direction update, one component: p <- z + beta p.  One spelling for every kernel that forms it
// (k_pcg_dir, k_dir_p2p, the fused small-graph kernel's on-the-fly form), so the rounding is the same everywhere.
GS_HD double dir_value(double z, double beta, double p) { return fma(beta, p, z); }

// p: owned rows; pg: ghost rows (local column c >= n_own reads pg + 3 (c - n_own)).
// DIR: the direction is not stored yet: p(c) = z(c) + beta * pold(c) is formed on the fly for every column the
// row touches (ghost columns from zg / pg), and the row's own new p goes to pnew.
template <bool DIR>
GS_HD double spmv_row_t(const SpmvView& a, const double* p, const double* pg, const double* z, const double* zg, double beta,
                        double* pnew, double* y, int r, double lambda) {
  const int sl = r / kSlice, lane = r % kSlice;
  const int base = a.rs.base[sl] + lane, w = a.rs.width[sl];
  double y0 = 0.0, y1 = 0.0, y2 = 0.0;
  const size_t n = (size_t)a.n_own, rr = (size_t)r;
  double d[6], l[3] = {lambda, lambda, lambda};
#if GS_SPMV_PRELOAD
#pragma unroll
  for (int k = 0; k < 6; ++k) d[k] = ldg(a.Hd + (size_t)k * n + rr);
#endif
  // w is uniform over the slice; the next slot's reference is fetched while this slot's block and
  // p are in flight, so a slot costs one dependent round trip
  int2 nref = w > 0 ? ldg(a.refs + base) : make_int2(-1, 0);
  for (int s = 0; s < w; ++s) {
    const int2 ref = nref;
    if (s + 1 < w) nref = ldg(a.refs + base + (s + 1) * kSlice);
    if (ref.x >= 0) {
      double h[9];
      load_block_as(a.Ho, a.hstride, ref.y >> 1, ref.y & 1, h);
      double px, py, pz;
      if (DIR) {
        const bool ghost = ref.x >= a.n_own;
        const size_t o = 3 * (size_t)(ghost ? ref.x - a.n_own : ref.x);
        const double* zc = (ghost ? zg : z) + o;
        const double* pc = (ghost ? pg : p) + o;
        px = dir_value(zc[0], beta, pc[0]);
        py = dir_value(zc[1], beta, pc[1]);
        pz = dir_value(zc[2], beta, pc[2]);
      } else {
        const double* pc = ref.x < a.n_own ? p + 3 * (size_t)ref.x : pg + 3 * (size_t)(ref.x - a.n_own);
        px = pc[0];
        py = pc[1];
        pz = pc[2];
      }
      spmv_slot(h, px, py, pz, y0, y1, y2);
    }
  }
#if !GS_SPMV_PRELOAD
#pragma unroll
  for (int k = 0; k < 6; ++k) d[k] = ldg(a.Hd + (size_t)k * n + rr);
#endif
  if (a.dscale) {
    l[0] *= a.dscale[3 * rr];
    l[1] *= a.dscale[3 * rr + 1];
    l[2] *= a.dscale[3 * rr + 2];
  }
  double q0 = p[3 * rr], q1 = p[3 * rr + 1], q2 = p[3 * rr + 2];
  if (DIR) {
    q0 = dir_value(z[3 * rr], beta, q0);
    q1 = dir_value(z[3 * rr + 1], beta, q1);
    q2 = dir_value(z[3 * rr + 2], beta, q2);
    pnew[3 * rr] = q0;
    pnew[3 * rr + 1] = q1;
    pnew[3 * rr + 2] = q2;
  }
  return spmv_finish(d, l, q0, q1, q2, y0, y1, y2, y, rr);
}

GS_HD double spmv_row(const SpmvView& a, const double* p, const double* pg, double* y, int r, double lambda) {
  return spmv_row_t<false>(a, p, pg, nullptr, nullptr, 0.0, nullptr, y, r, lambda);
}

// ------------------------------------------------------------------------------------------------
// Block-Jacobi preconditioner block and the PCG vector updates of row r
// ------------------------------------------------------------------------------------------------
GS_HD void minv_row(const double* Hd, const double* dscale, double* Minv, int n_own, int r, double lambda) {
  const size_t n = (size_t)n_own, v = (size_t)r;
  Sym3 A{Hd[v], Hd[n + v], Hd[2 * n + v], Hd[3 * n + v], Hd[4 * n + v], Hd[5 * n + v]};
  if (dscale) {
    A.a00 += lambda * dscale[3 * v];
    A.a11 += lambda * dscale[3 * v + 1];
    A.a22 += lambda * dscale[3 * v + 2];
  } else {
    A.a00 += lambda;
    A.a11 += lambda;
    A.a22 += lambda;
  }
  const Sym3 R = inverse_spd(A);
  Minv[v] = R.a00;
  Minv[n + v] = R.a01;
  Minv[2 * n + v] = R.a02;
  Minv[3 * n + v] = R.a11;
  Minv[4 * n + v] = R.a12;
  Minv[5 * n + v] = R.a22;
}

GS_HD void apply_minv(const double* Minv, size_t n, size_t v, double r0, double r1, double r2, double& z0, double& z1,
                      double& z2) {
  const double m00 = ldg(Minv + v), m01 = ldg(Minv + n + v), m02 = ldg(Minv + 2 * n + v);
  const double m11 = ldg(Minv + 3 * n + v), m12 = ldg(Minv + 4 * n + v), m22 = ldg(Minv + 5 * n + v);
  z0 = (m00 * r0 + m01 * r1) + m02 * r2;
  z1 = (m01 * r0 + m11 * r1) + m12 * r2;
  z2 = (m02 * r0 + m12 * r1) + m22 * r2;
}

struct PcgView {
  const double* b;     // gradient: rhs g = -b
  const double* Minv;  // 6 planes of n_own
  double* x;
  double* r;
  double* z;
  double* p;
  const double* Ap;
  int n_own;
};

// x = 0 ; r = -b ; z = Minv r ; p = z ; out: r.z, r.r
GS_HD void pcg_init_row(const PcgView& a, int v, double& rz, double& rr) {
  const size_t o = 3 * (size_t)v;
  const double r0 = -a.b[o], r1 = -a.b[o + 1], r2 = -a.b[o + 2];
  double z0, z1, z2;
  apply_minv(a.Minv, (size_t)a.n_own, (size_t)v, r0, r1, r2, z0, z1, z2);
  a.x[o] = 0.0; a.x[o + 1] = 0.0; a.x[o + 2] = 0.0;
  a.r[o] = r0;  a.r[o + 1] = r1;  a.r[o + 2] = r2;
  a.z[o] = z0;  a.z[o + 1] = z1;  a.z[o + 2] = z2;
  a.p[o] = z0;  a.p[o + 1] = z1;  a.p[o + 2] = z2;
  rz = (r0 * z0 + r1 * z1) + r2 * z2;
  rr = (r0 * r0 + r1 * r1) + r2 * r2;
}

// x += alpha p ; r -= alpha Ap ; z = Minv r ; out: r.z, r.r
// In two halves so that a kernel which has to wait for alpha (peer-memory path) can have the row's
// loads in flight while it waits; pcg_update_row is the two halves back to back (same arithmetic).
struct UpdateRow {
  double p0, p1, p2, q0, q1, q2, x0, x1, x2, r0, r1, r2;
  double m00, m01, m02, m11, m12, m22;
};

GS_HD void pcg_update_load(const PcgView& a, int v, UpdateRow& u) {
  const size_t o = 3 * (size_t)v, n = (size_t)a.n_own, w = (size_t)v;
  u.p0 = a.p[o]; u.p1 = a.p[o + 1]; u.p2 = a.p[o + 2];
  u.q0 = a.Ap[o]; u.q1 = a.Ap[o + 1]; u.q2 = a.Ap[o + 2];
  u.x0 = a.x[o]; u.x1 = a.x[o + 1]; u.x2 = a.x[o + 2];
  u.r0 = a.r[o]; u.r1 = a.r[o + 1]; u.r2 = a.r[o + 2];
  u.m00 = ldg(a.Minv + w); u.m01 = ldg(a.Minv + n + w); u.m02 = ldg(a.Minv + 2 * n + w);
  u.m11 = ldg(a.Minv + 3 * n + w); u.m12 = ldg(a.Minv + 4 * n + w); u.m22 = ldg(a.Minv + 5 * n + w);
}

GS_HD void pcg_update_apply(const PcgView& a, int v, const UpdateRow& u, double alpha, double& rz, double& rr) {
  const size_t o = 3 * (size_t)v;
  a.x[o] = u.x0 + alpha * u.p0;
  a.x[o + 1] = u.x1 + alpha * u.p1;
  a.x[o + 2] = u.x2 + alpha * u.p2;
  const double r0 = u.r0 - alpha * u.q0, r1 = u.r1 - alpha * u.q1, r2 = u.r2 - alpha * u.q2;
  const double z0 = (u.m00 * r0 + u.m01 * r1) + u.m02 * r2;
  const double z1 = (u.m01 * r0 + u.m11 * r1) + u.m12 * r2;
  const double z2 = (u.m02 * r0 + u.m12 * r1) + u.m22 * r2;
  a.r[o] = r0; a.r[o + 1] = r1; a.r[o + 2] = r2;
  a.z[o] = z0; a.z[o + 1] = z1; a.z[o + 2] = z2;
  rz = (r0 * z0 + r1 * z1) + r2 * z2;
  rr = (r0 * r0 + r1 * r1) + r2 * r2;
}

GS_HD void pcg_update_row(const PcgView& a, int v, double alpha, double& rz, double& rr) {
  UpdateRow u;
  pcg_update_load(a, v, u);
  pcg_update_apply(a, v, u, alpha, rz, rr);
}

// p = z + beta p
GS_HD void pcg_dir_row(const PcgView& a, int v, double beta) {
  const size_t o = 3 * (size_t)v;
  a.p[o] = dir_value(a.z[o], beta, a.p[o]);
  a.p[o + 1] = dir_value(a.z[o + 1], beta, a.p[o + 1]);
  a.p[o + 2] = dir_value(a.z[o + 2], beta, a.p[o + 2]);
}

GS_HD double dot_row(const double* u, const double* w, int v) {
  const size_t o = 3 * (size_t)v;
  return (u[o] * w[o] + u[o + 1] * w[o + 1]) + u[o + 2] * w[o + 2];
}

}  // namespace gs
