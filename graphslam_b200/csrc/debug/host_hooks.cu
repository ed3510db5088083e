// TEST-ONLY host entry points (include/graphslam_cuda_debug.h), built into libgraphslam_cuda_debug.so together with
// the CPU emulator (emu.cu) — never into the product library.  They call the product's own headers (structure build,
// g2o reader/writer, per-factor inline functions), so the CPU tests exercise the code the product compiles.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../../include/graphslam_cuda_debug.h"
#include "../g2o.hpp"
#include "../host_util.hpp"
#include "../rows.cuh"

using gs::spd_information;

extern "C" {

gs_status gs_host_pattern(int32_t n, int32_t e, const int32_t* ei, const int32_t* ej, int32_t* rowptr, int32_t* colidx,
                          size_t* nnzb, int32_t* perm) try {
  if (n <= 0 || e < 0 || (e && (!ei || !ej))) return GS_EINVAL;
  std::vector<int32_t> rp, ci;
  gs::global_pattern(n, e, ei, ej, rp, ci);
  if (nnzb) *nnzb = ci.size();
  if (rowptr) std::memcpy(rowptr, rp.data(), sizeof(int32_t) * rp.size());
  if (colidx) std::memcpy(colidx, ci.data(), sizeof(int32_t) * ci.size());
  if (perm) {
    std::vector<int> pm;
    gs::canonical_order(n, e, ei, ej, pm);
    std::memcpy(perm, pm.data(), sizeof(int32_t) * e);
  }
  return GS_OK;
} catch (...) {
  return GS_ENOMEM;
}

// the aggregates of the two-level preconditioner's coarse space (structure.hpp: build_aggregation, as prepare() calls it):
// counts = {aggregates, colours, target size}; agg_of (n), color (n_agg, sized from a first call), seg_agg (9) may be NULL
gs_status gs_host_aggregation(int32_t n, int32_t e, const int32_t* ei, const int32_t* ej, int32_t target, int32_t counts[3],
                              int32_t* agg_of, int32_t* color, int32_t* seg_agg) try {
  if (n <= 0 || e < 0 || (e && (!ei || !ej)) || !counts) return GS_EINVAL;
  std::vector<int32_t> rp, ci;
  gs::global_pattern(n, e, ei, ej, rp, ci);
  const int t = target > 0 ? target : std::max(48, (int)std::ceil(std::sqrt((double)n)));
  gs::Aggregation A;
  gs::build_aggregation(n, rp, ci, std::max(t, (n + 4095) / 4096), A);
  counts[0] = A.n_agg;
  counts[1] = A.n_colors;
  counts[2] = A.target;
  if (agg_of) std::memcpy(agg_of, A.agg_of.data(), sizeof(int32_t) * (size_t)n);
  if (color) std::memcpy(color, A.color.data(), sizeof(int32_t) * (size_t)A.n_agg);
  if (seg_agg) std::memcpy(seg_agg, A.seg_agg, sizeof(int32_t) * (gs::kSegments + 1));
  return GS_OK;
} catch (...) {
  return GS_ENOMEM;
}

// the g2o reader/writer the product uses, without a graph handle: parse `in_path`, report the counts
// {vertices, edges, priors, fixed, skipped}, optionally copy the parsed arrays out and/or write the
// parsed content back to `out_path`
gs_status gs_host_g2o(const char* in_path, const char* out_path, size_t counts[5], uint64_t* vkey, double* vxyt,
                      uint64_t* e1, uint64_t* e2, double* ez, double* einfo, uint64_t* pkey, double* pxyt, double* pinfo) try {
  if (!in_path) return GS_EINVAL;
  gs::G2oData d;
  if (!gs::read_g2o(in_path, d).empty()) return GS_EINVAL;
  if (counts) {
    counts[0] = d.vkey.size();
    counts[1] = d.e1.size();
    counts[2] = d.pkey.size();
    counts[3] = d.fixed.size();
    counts[4] = d.skipped;
  }
  auto out = [](auto* dst, const auto& v) {
    if (dst && !v.empty()) std::memcpy(dst, v.data(), sizeof(v[0]) * v.size());
  };
  out(vkey, d.vkey);
  out(vxyt, d.vxyt);
  out(e1, d.e1);
  out(e2, d.e2);
  out(ez, d.ez);
  out(einfo, d.einfo);
  out(pkey, d.pkey);
  out(pxyt, d.pxyt);
  out(pinfo, d.pinfo);
  if (out_path && !gs::write_g2o(out_path, d).empty()) return GS_EINVAL;
  return GS_OK;
} catch (...) {
  return GS_ENOMEM;
}

// one factor, host evaluation of the same inline functions the kernels use
gs_status gs_host_between(const double pi[3], const double pj[3], const double z[3], const double cov9[9], int mode,
                          double e[3], double Ai[9], double Aj[9], double Hii[9], double Hij[9], double Hjj[9],
                          double bi[3], double bj[3]) try {
  double om[6];
  if (!spd_information(cov9, om)) return GS_ENOTSPD;
  const double4 a = make_double4(pi[0], pi[1], std::cos(pi[2]), std::sin(pi[2]));
  const double4 b = make_double4(pj[0], pj[1], std::cos(pj[2]), std::sin(pj[2]));
  const double2 zt = make_double2(z[0], z[1]), zr = make_double2(std::cos(z[2]), std::sin(z[2]));
  gs::BetweenLin L;
  if (mode == 0) gs::between_linearize<0>(a, b, zt, zr, L);
  else gs::between_linearize<1>(a, b, zt, zr, L);
  gs::Sym3 O{om[0], om[1], om[2], om[3], om[4], om[5]};
  double dg[6], bv[3];
  gs::diag_pieces(O, L.Ai, L.e, dg, bv);
  const double fi[9] = {dg[0], dg[1], dg[2], dg[1], dg[3], dg[4], dg[2], dg[4], dg[5]};
  std::memcpy(Hii, fi, sizeof fi);
  std::memcpy(bi, bv, sizeof bv);
  gs::off_block(O, L.Ai, L.Aj, Hij);
  gs::diag_pieces(O, L.Aj, L.e, dg, bv);
  const double fj[9] = {dg[0], dg[1], dg[2], dg[1], dg[3], dg[4], dg[2], dg[4], dg[5]};
  std::memcpy(Hjj, fj, sizeof fj);
  std::memcpy(bj, bv, sizeof bv);
  std::memcpy(e, L.e, sizeof L.e);
  std::memcpy(Ai, L.Ai.m, sizeof L.Ai.m);
  std::memcpy(Aj, L.Aj.m, sizeof L.Aj.m);
  return GS_OK;
} catch (...) {
  return GS_ENOMEM;
}

}  // extern "C"
