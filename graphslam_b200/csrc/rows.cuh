// Per-row bodies of the Gauss-Newton / Levenberg-Marquardt kernels.  One thread owns one block row
// (one vertex): it walks the row's SELL-32 list, so the 32 lanes of a warp read 32 consecutive
// words per slot (coalesced), every output has exactly one writer (no atomics) and every sum has
// one fixed order.  The functions are __host__ __device__: the kernels in kernels.cuh call them per
// thread, and the test-only emulator (emu.cu) calls the very same code per row on the CPU.
#pragma once
#include <cstdint>
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

struct FactorView {
  const int2* ij;       // local vertex indices of id_1, id_2
  const double2* zt;    // measured translation
  const double2* zr;    // (cos, sin) of the measured rotation
  const double2* o01;   // information, upper triangle 00 01 | 02 11 | 12 22
  const double2* o23;
  const double2* o45;
};

struct PriorView {
  const int* row;        // local row
  const double4* pose;   // (x, y, cos, sin)
  const double* om;      // 6 per prior
};

// ------------------------------------------------------------------------------------------------
// K1+K2: linearise every factor side of row r and reduce it into the row's diagonal block, its
// gradient, and the off-diagonal blocks this row is the writer of.
// ------------------------------------------------------------------------------------------------
struct AsmView {
  SellView cs;
  const uint32_t* c_code;
  const int* c_blk;
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
#pragma unroll
  for (int k = 0; k < 9; ++k) Ho[(size_t)k * hstride + pos] = M[k];
}

// returns the row's chi2 share (factors it is the counting side of, and its priors)
template <int MODE>
GS_HD double assemble_row(const AsmView& a, int r) {
  const int sl = r / kSlice, lane = r % kSlice;
  const int base = a.cs.base[sl] + lane, w = a.cs.width[sl];
  double dg[6] = {0, 0, 0, 0, 0, 0}, bv[3] = {0, 0, 0}, acc[9];
  double chi2 = 0.0;
  int open = -1;
  for (int s = 0; s < w; ++s) {
    const int c = base + s * kSlice;
    const uint32_t code = ldg(a.c_code + c);
    if (code == kPadCode) break;
    const int cb = ldg(a.c_blk + c);
    const int kind = code & 3u;
    const int idx = (int)(code >> 3);
    double d6[6], b3[3];
    if (kind < 2) {
      const int2 ij = ldg(a.f.ij + idx);
      const double2 zt = ldg(a.f.zt + idx), zr = ldg(a.f.zr + idx);
      const double2 o01 = ldg(a.f.o01 + idx), o23 = ldg(a.f.o23 + idx), o45 = ldg(a.f.o45 + idx);
      const double4 pi = ldg(a.pose + ij.x), pj = ldg(a.pose + ij.y);
      BetweenLin L;
      between_linearize<MODE>(pi, pj, zt, zr, L);
      const Sym3 O{o01.x, o01.y, o23.x, o23.y, o45.x, o45.y};
      const bool row_is_i = (kind == 0);
      diag_pieces(O, row_is_i ? L.Ai : L.Aj, L.e, d6, b3);
      if (code & 4u) chi2 += quad_form(O, L.e);
      if (cb >= 0) {
        // row is the hi endpoint of the pair (cb & 2): lo Jacobian is the other vertex's
        const bool lo_is_i = row_is_i != ((cb & 2) != 0);
        double M[9];
        off_block(O, lo_is_i ? L.Ai : L.Aj, lo_is_i ? L.Aj : L.Ai, M);
        if (cb & 1) {
          if (open >= 0) store_block(a.Ho, a.hstride, open, acc);
          open = cb >> 2;
#pragma unroll
          for (int k = 0; k < 9; ++k) acc[k] = M[k];
        } else {
#pragma unroll
          for (int k = 0; k < 9; ++k) acc[k] += M[k];
        }
      }
    } else {
      const double4 pp = ldg(a.pr.pose + idx);
      const double4 px = ldg(a.pose + r);
      const double* om = a.pr.om + 6 * idx;
      const Sym3 O{om[0], om[1], om[2], om[3], om[4], om[5]};
      double e[3];
      Mat3 A;
      prior_linearize<MODE>(px, pp, e, A);
      diag_pieces(O, A, e, d6, b3);
      chi2 += quad_form(O, e);
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) dg[k] += d6[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) bv[k] += b3[k];
  }
  if (open >= 0) store_block(a.Ho, a.hstride, open, acc);
#pragma unroll
  for (int k = 0; k < 6; ++k) a.Hd[(size_t)k * a.n_own + r] = dg[k];
#pragma unroll
  for (int k = 0; k < 3; ++k) a.b[3 * (size_t)r + k] = bv[k];
  return chi2;
}

// graph.error contribution of row r (same counting rule as assemble_row), at `pose`
GS_HD double chi2_row(const AsmView& a, int r) {
  const int sl = r / kSlice, lane = r % kSlice;
  const int base = a.cs.base[sl] + lane, w = a.cs.width[sl];
  double chi2 = 0.0;
  for (int s = 0; s < w; ++s) {
    const int c = base + s * kSlice;
    const uint32_t code = ldg(a.c_code + c);
    if (code == kPadCode) break;
    if (!(code & 4u)) continue;
    const int kind = code & 3u;
    const int idx = (int)(code >> 3);
    double e[3];
    if (kind < 2) {
      const int2 ij = ldg(a.f.ij + idx);
      const double2 zt = ldg(a.f.zt + idx), zr = ldg(a.f.zr + idx);
      const double2 o01 = ldg(a.f.o01 + idx), o23 = ldg(a.f.o23 + idx), o45 = ldg(a.f.o45 + idx);
      between_error(ldg(a.pose + ij.x), ldg(a.pose + ij.y), zt, zr, e);
      const Sym3 O{o01.x, o01.y, o23.x, o23.y, o45.x, o45.y};
      chi2 += quad_form(O, e);
    } else {
      const double* om = a.pr.om + 6 * idx;
      const Sym3 O{om[0], om[1], om[2], om[3], om[4], om[5]};
      Mat3 A;
      prior_linearize<0>(ldg(a.pose + r), ldg(a.pr.pose + idx), e, A);
      chi2 += quad_form(O, e);
    }
  }
  return chi2;
}

// ------------------------------------------------------------------------------------------------
// K3: y_r = sum_c H_rc p_c (global column order) + H_rr p_r + damp p_r ; returns p_r . y_r
// ------------------------------------------------------------------------------------------------
struct SpmvView {
  SellView rs;
  const int* r_col;
  const int* r_blk;
  const double* Ho;
  size_t hstride;
  const double* Hd;      // 6 planes of n_own
  const double* dscale;  // nullptr: lambda*I ; else lambda*diag(dscale), 3 per row
  int n_own;
};

GS_HD double spmv_row(const SpmvView& a, const double* p, double* y, int r, double lambda) {
  const int sl = r / kSlice, lane = r % kSlice;
  const int base = a.rs.base[sl] + lane, w = a.rs.width[sl];
  double y0 = 0.0, y1 = 0.0, y2 = 0.0;
  for (int s = 0; s < w; ++s) {
    const int q = base + s * kSlice;
    const int col = ldg(a.r_col + q);
    if (col < 0) break;
    const int rb = ldg(a.r_blk + q);
    const int blk = rb >> 1;
    const double* h = a.Ho + blk;
    const double h0 = ldg(h), h1 = ldg(h + a.hstride), h2 = ldg(h + 2 * a.hstride);
    const double h3 = ldg(h + 3 * a.hstride), h4 = ldg(h + 4 * a.hstride), h5 = ldg(h + 5 * a.hstride);
    const double h6 = ldg(h + 6 * a.hstride), h7 = ldg(h + 7 * a.hstride), h8 = ldg(h + 8 * a.hstride);
    const double px = p[3 * (size_t)col], py = p[3 * (size_t)col + 1], pz = p[3 * (size_t)col + 2];
    const bool tr = rb & 1;
    const double a01 = tr ? h3 : h1, a02 = tr ? h6 : h2, a10 = tr ? h1 : h3;
    const double a12 = tr ? h7 : h5, a20 = tr ? h2 : h6, a21 = tr ? h5 : h7;
    y0 += (h0 * px + a01 * py) + a02 * pz;
    y1 += (a10 * px + h4 * py) + a12 * pz;
    y2 += (a20 * px + a21 * py) + h8 * pz;
  }
  const size_t n = (size_t)a.n_own, rr = (size_t)r;
  const double d00 = ldg(a.Hd + rr), d01 = ldg(a.Hd + n + rr), d02 = ldg(a.Hd + 2 * n + rr);
  const double d11 = ldg(a.Hd + 3 * n + rr), d12 = ldg(a.Hd + 4 * n + rr), d22 = ldg(a.Hd + 5 * n + rr);
  const double px = p[3 * rr], py = p[3 * rr + 1], pz = p[3 * rr + 2];
  double l0 = lambda, l1 = lambda, l2 = lambda;
  if (a.dscale) {
    l0 *= a.dscale[3 * rr];
    l1 *= a.dscale[3 * rr + 1];
    l2 *= a.dscale[3 * rr + 2];
  }
  y0 = (y0 + ((d00 * px + d01 * py) + d02 * pz)) + l0 * px;
  y1 = (y1 + ((d01 * px + d11 * py) + d12 * pz)) + l1 * py;
  y2 = (y2 + ((d02 * px + d12 * py) + d22 * pz)) + l2 * pz;
  y[3 * rr] = y0;
  y[3 * rr + 1] = y1;
  y[3 * rr + 2] = y2;
  return (px * y0 + py * y1) + pz * y2;
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
GS_HD void pcg_update_row(const PcgView& a, int v, double alpha, double& rz, double& rr) {
  const size_t o = 3 * (size_t)v;
  const double p0 = a.p[o], p1 = a.p[o + 1], p2 = a.p[o + 2];
  const double q0 = a.Ap[o], q1 = a.Ap[o + 1], q2 = a.Ap[o + 2];
  a.x[o] += alpha * p0;
  a.x[o + 1] += alpha * p1;
  a.x[o + 2] += alpha * p2;
  const double r0 = a.r[o] - alpha * q0, r1 = a.r[o + 1] - alpha * q1, r2 = a.r[o + 2] - alpha * q2;
  double z0, z1, z2;
  apply_minv(a.Minv, (size_t)a.n_own, (size_t)v, r0, r1, r2, z0, z1, z2);
  a.r[o] = r0; a.r[o + 1] = r1; a.r[o + 2] = r2;
  a.z[o] = z0; a.z[o + 1] = z1; a.z[o + 2] = z2;
  rz = (r0 * z0 + r1 * z1) + r2 * z2;
  rr = (r0 * r0 + r1 * r1) + r2 * r2;
}

// p = z + beta p
GS_HD void pcg_dir_row(const PcgView& a, int v, double beta) {
  const size_t o = 3 * (size_t)v;
  a.p[o] = a.z[o] + beta * a.p[o];
  a.p[o + 1] = a.z[o + 1] + beta * a.p[o + 1];
  a.p[o + 2] = a.z[o + 2] + beta * a.p[o + 2];
}

GS_HD double dot_row(const double* u, const double* w, int v) {
  const size_t o = 3 * (size_t)v;
  return (u[o] * w[o] + u[o + 1] * w[o + 1]) + u[o + 2] * w[o + 2];
}

}  // namespace gs
