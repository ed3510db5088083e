// Host-side checks shared by the library and the test-only debug library.
#pragma once
#include <algorithm>
#include <cmath>
#include <thread>
#include <vector>
#include "host_pool.hpp"
#include "se2.cuh"

namespace gs {

// fn(begin, end) over [0, n) on up to 16 host threads of the process-wide pool (per-pose conversions between the caller's
// (x, y, theta) and the device's (x, y, cos, sin): a million sin/cos/atan2 on one thread is 20-50 ms, which a 140-ms LM
// step notices)
template <typename F>
inline void parallel_for(size_t n, F fn, size_t min_chunk = 32768) {
  size_t nt = std::min<size_t>(std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), 16), (n + min_chunk - 1) / min_chunk);
  if (nt <= 1) {
    fn((size_t)0, n);
    return;
  }
  const size_t chunk = (n + nt - 1) / nt;
  HostPool::get().run((int)nt, [&](int t) {
    const size_t b = (size_t)t * chunk, e = std::min(n, b + chunk);
    if (b < e) fn(b, e);
  });
}

// noiseModel::Gaussian::Covariance(Q): symmetrise, Cholesky test (GTSAM throws on a non-SPD Q), invert -> information
inline bool spd_information(const double* cov9, double* om6) {
  // symmetrise, Cholesky test, invert via adjugate
  const double a00 = cov9[0], a11 = cov9[4], a22 = cov9[8];
  const double a01 = 0.5 * (cov9[1] + cov9[3]), a02 = 0.5 * (cov9[2] + cov9[6]), a12 = 0.5 * (cov9[5] + cov9[7]);
  if (!(a00 > 0.0)) return false;
  const double l00 = std::sqrt(a00), l10 = a01 / l00, l20 = a02 / l00;
  const double d1 = a11 - l10 * l10;
  if (!(d1 > 0.0)) return false;
  const double l11 = std::sqrt(d1), l21 = (a12 - l20 * l10) / l11;
  const double d2 = a22 - l20 * l20 - l21 * l21;
  if (!(d2 > 0.0) || !std::isfinite(d2)) return false;
  gs::Sym3 q{a00, a01, a02, a11, a12, a22};
  const gs::Sym3 r = gs::inverse_spd(q);
  om6[0] = r.a00; om6[1] = r.a01; om6[2] = r.a02; om6[3] = r.a11; om6[4] = r.a12; om6[5] = r.a22;
  for (int k = 0; k < 6; ++k)
    if (!std::isfinite(om6[k])) return false;
  return true;
}

// noiseModel::Gaussian::Information(I): information (inverse covariance) given directly, upper triangle I11 I12 I13 I22 I23 I33: SPD test only
inline bool spd_upper(const double* o) {
  for (int k = 0; k < 6; ++k)
    if (!std::isfinite(o[k])) return false;
  if (!(o[0] > 0.0)) return false;
  const double l00 = std::sqrt(o[0]), l10 = o[1] / l00, l20 = o[2] / l00;
  const double d1 = o[3] - l10 * l10;
  if (!(d1 > 0.0)) return false;
  const double l11 = std::sqrt(d1), l21 = (o[4] - l20 * l10) / l11;
  const double d2 = o[5] - l20 * l20 - l21 * l21;
  return d2 > 0.0;
}


}  // namespace gs
