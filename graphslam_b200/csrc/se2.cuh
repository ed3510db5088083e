// SE(2) arithmetic shared by every kernel (FP64, no fast-math).
//
// Restates the GTSAM 4.0.x Pose2 semantics that the reference selects through
// BetweenFactor<Pose2> / PriorFactor<Pose2> (reference src/graph/src/graph.cpp:57,87-92,106-111):
//   * a pose is stored as (x, y, cos th, sin th) like gtsam::Pose2 (translation + Rot2(c,s));
//     32 B, so one gather is one DRAM/L2 sector.
//   * between(p1,p2) = p1^-1 o p2 with H1 = [[-c,-s,dt1],[s,-c,dt2],[0,0,-1]], H2 = I
//   * BetweenFactor error e = Local(z^-1 o between) ; Local(p) = (x, y, theta), theta = atan2(s,c)
//   * retract(p, v) = p o Pose2(v)  (body-frame update)
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#define GS_HD __host__ __device__ __forceinline__

namespace gs {

// gtsam::Rot2::normalize(): rescale (c, s) only when c^2 + s^2 is off 1 by more than 1e-10 — which the product of
// two unit rotations never is, so in the kernels this is one compare and an untaken branch, not a sqrt and a
// division in every factor's dependency chain (Rot2::atan2 / Rot2::fromCosSin both go through it).
GS_HD void rot2_normalize(double& c, double& s) {
  const double scale = c * c + s * s;
  if (fabs(scale - 1.0) > 1e-10) {
    const double inv = 1.0 / sqrt(scale);
    c *= inv;
    s *= inv;
  }
}

struct Sym3 {  // symmetric 3x3: 00 01 02 11 12 22
  double a00, a01, a02, a11, a12, a22;
};

struct Mat3 {  // row-major
  double m[9];
};

// Residual and Jacobians of one BetweenFactor.
struct BetweenLin {
  double e[3];
  Mat3 Ai, Aj;
};

template <int MODE>
GS_HD void between_linearize(const double4 pi, const double4 pj, const double2 zt, const double2 zr, BetweenLin& L) {
  const double ci = pi.z, si = pi.w, cj = pj.z, sj = pj.w;
  const double dx = pj.x - pi.x, dy = pj.y - pi.y;
  const double hx = ci * dx + si * dy;
  const double hy = -si * dx + ci * dy;
  double c = ci * cj + si * sj;
  double s = -si * cj + ci * sj;
  rot2_normalize(c, s);  // Rot2::atan2(s, c)
  const double cm = zr.x, sm = zr.y;
  const double rx = hx - zt.x, ry = hy - zt.y;
  L.e[0] = cm * rx + sm * ry;
  L.e[1] = -sm * rx + cm * ry;
  const double se = s * cm - c * sm, ce = c * cm + s * sm;
  L.e[2] = atan2(se, ce);
  const double dt1 = -sj * dx + cj * dy;
  const double dt2 = -cj * dx - sj * dy;
  if (MODE == 0) {
    L.Ai.m[0] = -c;  L.Ai.m[1] = -s;  L.Ai.m[2] = dt1;
    L.Ai.m[3] = s;   L.Ai.m[4] = -c;  L.Ai.m[5] = dt2;
    L.Ai.m[6] = 0.0; L.Ai.m[7] = 0.0; L.Ai.m[8] = -1.0;
    L.Aj.m[0] = 1.0; L.Aj.m[1] = 0.0; L.Aj.m[2] = 0.0;
    L.Aj.m[3] = 0.0; L.Aj.m[4] = 1.0; L.Aj.m[5] = 0.0;
    L.Aj.m[6] = 0.0; L.Aj.m[7] = 0.0; L.Aj.m[8] = 1.0;
  } else {
    // exact: D * H with D = blkdiag(R(e_theta), 1); (ce, se) is already (cos, sin) of e_theta
    const double n = 1.0 / sqrt(ce * ce + se * se);
    const double cE = ce * n, sE = se * n;
    L.Ai.m[0] = cE * (-c) - sE * s;   L.Ai.m[1] = cE * (-s) - sE * (-c);  L.Ai.m[2] = cE * dt1 - sE * dt2;
    L.Ai.m[3] = sE * (-c) + cE * s;   L.Ai.m[4] = sE * (-s) + cE * (-c);  L.Ai.m[5] = sE * dt1 + cE * dt2;
    L.Ai.m[6] = 0.0;                  L.Ai.m[7] = 0.0;                    L.Ai.m[8] = -1.0;
    L.Aj.m[0] = cE;  L.Aj.m[1] = -sE; L.Aj.m[2] = 0.0;
    L.Aj.m[3] = sE;  L.Aj.m[4] = cE;  L.Aj.m[5] = 0.0;
    L.Aj.m[6] = 0.0; L.Aj.m[7] = 0.0; L.Aj.m[8] = 1.0;
  }
}

// Residual only (chi2 evaluation).
GS_HD void between_error(const double4 pi, const double4 pj, const double2 zt, const double2 zr, double e[3]) {
  const double ci = pi.z, si = pi.w, cj = pj.z, sj = pj.w;
  const double dx = pj.x - pi.x, dy = pj.y - pi.y;
  const double hx = ci * dx + si * dy;
  const double hy = -si * dx + ci * dy;
  double c = ci * cj + si * sj;
  double s = -si * cj + ci * sj;
  rot2_normalize(c, s);
  const double cm = zr.x, sm = zr.y;
  const double rx = hx - zt.x, ry = hy - zt.y;
  e[0] = cm * rx + sm * ry;
  e[1] = -sm * rx + cm * ry;
  e[2] = atan2(s * cm - c * sm, c * cm + s * sm);
}

// PriorFactor<Pose2> (GTSAM 4.0): e = -Local(x^-1 o prior), H = I (exact mode: d e / d xi).
template <int MODE>
GS_HD void prior_linearize(const double4 p, const double4 prior, double e[3], Mat3& A) {
  const double c = p.z, s = p.w;
  const double dx = prior.x - p.x, dy = prior.y - p.y;
  e[0] = -(c * dx + s * dy);
  e[1] = -(-s * dx + c * dy);
  // theta_prior - theta_x through (c,s): sin = sp*c - cp*s ; cos = cp*c + sp*s
  e[2] = -atan2(prior.w * c - prior.z * s, prior.z * c + prior.w * s);
  A.m[0] = 1.0; A.m[1] = 0.0; A.m[2] = 0.0;
  A.m[3] = 0.0; A.m[4] = 1.0; A.m[5] = 0.0;
  A.m[6] = 0.0; A.m[7] = 0.0; A.m[8] = 1.0;
  if (MODE != 0) {
    A.m[2] = e[1];
    A.m[5] = -e[0];
  }
}

GS_HD double quad_form(const Sym3& O, const double e[3]) {
  const double w0 = O.a00 * e[0] + O.a01 * e[1] + O.a02 * e[2];
  const double w1 = O.a01 * e[0] + O.a11 * e[1] + O.a12 * e[2];
  const double w2 = O.a02 * e[0] + O.a12 * e[1] + O.a22 * e[2];
  return e[0] * w0 + e[1] * w1 + e[2] * w2;
}

// One side of a factor: with Ar = Jacobian of the row vertex, Ac = Jacobian of the column vertex,
//   W = Omega*Ar ; diag = Ar^T W (sym) ; off = W^T Ac (= Ar^T Omega Ac) ; b = W^T e (= Ar^T Omega e)
GS_HD void normal_pieces(const Sym3& O, const Mat3& Ar, const Mat3& Ac, const double e[3], double off[9],
                         double diag[6], double bv[3]) {
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
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) off[3 * r + c] = W[0 + r] * Ac.m[0 + c] + W[3 + r] * Ac.m[3 + c] + W[6 + r] * Ac.m[6 + c];
    bv[r] = W[0 + r] * e[0] + W[3 + r] * e[1] + W[6 + r] * e[2];
  }
}

// Values::retract for Pose2: p o (dx, dy, dth).  The rotation product is ALWAYS renormalised here (so (c, s) cannot drift
// over thousands of accepted steps of a resident graph); gtsam::Rot2::fromCosSin renormalises only when c^2 + s^2 is
// off 1 by more than 1e-10, i.e. never for one product: the two differ by less than an ulp of theta = atan2(s, c),
// which is scale-invariant.  The C port and the numpy oracle normalise the same way.
GS_HD double4 retract(const double4 p, const double dx, const double dy, const double dth) {
  double sd, cd;
  sincos(dth, &sd, &cd);
  double4 q;
  q.x = p.x + p.z * dx - p.w * dy;
  q.y = p.y + p.w * dx + p.z * dy;
  double c = p.z * cd - p.w * sd;
  double s = p.w * cd + p.z * sd;
  const double inv = 1.0 / sqrt(c * c + s * s);
  q.z = c * inv;
  q.w = s * inv;
  return q;
}

// Inverse of a symmetric positive definite 3x3 (block-Jacobi preconditioner block).
GS_HD Sym3 inverse_spd(const Sym3& A) {
  const double c00 = A.a11 * A.a22 - A.a12 * A.a12;
  const double c01 = A.a02 * A.a12 - A.a01 * A.a22;
  const double c02 = A.a01 * A.a12 - A.a02 * A.a11;
  const double det = A.a00 * c00 + A.a01 * c01 + A.a02 * c02;
  const double id = 1.0 / det;
  Sym3 R;
  R.a00 = c00 * id;
  R.a01 = c01 * id;
  R.a02 = c02 * id;
  R.a11 = (A.a00 * A.a22 - A.a02 * A.a02) * id;
  R.a12 = (A.a01 * A.a02 - A.a00 * A.a12) * id;
  R.a22 = (A.a00 * A.a11 - A.a01 * A.a01) * id;
  return R;
}

}  // namespace gs
