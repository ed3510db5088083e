// g2o / TORO 2-D text wire format.  The reference keeps its graph only in
// process globals (src/graph/src/graph.cpp:5-10) and loses it on exit; this is the loader/saver that
// lets public SE(2) datasets (M3500, Intel, MIT) and checkpoints through the same C ABI.
//
//   VERTEX_SE2 id x y theta                                   (TORO: VERTEX2)
//   EDGE_SE2   i j dx dy dtheta I11 I12 I13 I22 I23 I33       (upper triangle of the information matrix)
//   EDGE2      i j dx dy dtheta I11 I12 I22 I33 I13 I23       (TORO's order)
//   EDGE_PRIOR_SE2 id x y theta I11 I12 I13 I22 I23 I33       (unary; how gs_save_g2o writes PriorFactor<Pose2>)
//   FIX id ...                                                (counted, not applied: the reference anchors the
//                                                              gauge with a PriorFactor, graph.cpp:57)
// '#' starts a comment.  Unknown tags are counted and skipped, malformed known lines are errors.
// Pure C++ (no CUDA): the CPU tests parse with exactly this code.
#pragma once
#include <locale.h>
#include <stdlib.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace gs {

struct G2oData {
  std::vector<uint64_t> vkey;
  std::vector<double> vxyt;              // 3 per vertex
  std::vector<uint64_t> e1, e2;
  std::vector<double> ez;                // 3 per edge
  std::vector<double> einfo;             // 6 per edge: I11 I12 I13 I22 I23 I33
  std::vector<uint64_t> pkey;
  std::vector<double> pxyt, pinfo;       // priors
  std::vector<uint64_t> fixed;           // FIX ids (reported only)
  size_t skipped = 0;                    // lines with an unknown tag
};

namespace g2o_detail {
// the "C" locale, whatever the host program has set with setlocale (a decimal comma would split "1.5" in two)
inline locale_t c_locale() {
  static locale_t loc = newlocale(LC_ALL_MASK, "C", (locale_t)0);
  return loc;
}
// parse `n` doubles from *p; advances *p; false when fewer are present
inline bool take_doubles(const char** p, int n, double* out) {
  for (int k = 0; k < n; ++k) {
    char* end = nullptr;
    out[k] = strtod_l(*p, &end, c_locale());
    if (end == *p) return false;
    *p = end;
  }
  return true;
}
inline bool take_key(const char** p, uint64_t* out) {
  while (**p == ' ' || **p == '\t') ++*p;
  if (**p == '-' ) return false;
  char* end = nullptr;
  *out = std::strtoull(*p, &end, 10);
  if (end == *p) return false;
  *p = end;
  return true;
}
}  // namespace g2o_detail

// returns "" on success, else a message naming the line
inline std::string read_g2o(const char* path, G2oData& d) {
  std::FILE* f = std::fopen(path, "r");
  if (!f) return std::string("cannot open ") + path;
  std::string err;
  std::vector<char> buf(1 << 16);
  size_t lineno = 0;
  while (std::fgets(buf.data(), (int)buf.size(), f)) {
    ++lineno;
    char* line = buf.data();
    if (char* hash = std::strchr(line, '#')) *hash = 0;
    while (*line == ' ' || *line == '\t') ++line;
    if (*line == 0 || *line == '\n' || *line == '\r') continue;
    char tag[64];
    int used = 0;
    if (std::sscanf(line, "%63s%n", tag, &used) != 1) continue;
    const char* p = line + used;
    auto bad = [&](const char* what) {
      err = std::string(path) + ":" + std::to_string(lineno) + ": malformed " + what + " line";
    };
    if (!std::strcmp(tag, "VERTEX_SE2") || !std::strcmp(tag, "VERTEX2")) {
      uint64_t id;
      double v[3];
      if (!g2o_detail::take_key(&p, &id) || !g2o_detail::take_doubles(&p, 3, v)) {
        bad(tag);
        break;
      }
      d.vkey.push_back(id);
      d.vxyt.insert(d.vxyt.end(), v, v + 3);
    } else if (!std::strcmp(tag, "EDGE_SE2") || !std::strcmp(tag, "EDGE2") || !std::strcmp(tag, "EDGE")) {
      uint64_t i, j;
      double v[9];
      if (!g2o_detail::take_key(&p, &i) || !g2o_detail::take_key(&p, &j) || !g2o_detail::take_doubles(&p, 9, v)) {
        bad(tag);
        break;
      }
      d.e1.push_back(i);
      d.e2.push_back(j);
      d.ez.insert(d.ez.end(), v, v + 3);
      if (!std::strcmp(tag, "EDGE_SE2")) {
        d.einfo.insert(d.einfo.end(), v + 3, v + 9);
      } else {  // TORO: I11 I12 I22 I33 I13 I23
        const double o[6] = {v[3], v[4], v[7], v[5], v[8], v[6]};
        d.einfo.insert(d.einfo.end(), o, o + 6);
      }
    } else if (!std::strcmp(tag, "EDGE_PRIOR_SE2")) {
      uint64_t id;
      double v[9];
      if (!g2o_detail::take_key(&p, &id) || !g2o_detail::take_doubles(&p, 9, v)) {
        bad(tag);
        break;
      }
      d.pkey.push_back(id);
      d.pxyt.insert(d.pxyt.end(), v, v + 3);
      d.pinfo.insert(d.pinfo.end(), v + 3, v + 9);
    } else if (!std::strcmp(tag, "FIX")) {
      uint64_t id;
      while (g2o_detail::take_key(&p, &id)) d.fixed.push_back(id);
    } else {
      d.skipped += 1;
    }
  }
  std::fclose(f);
  return err;
}

// %.17g round-trips every finite double exactly
inline std::string write_g2o(const char* path, const G2oData& d) {
  std::FILE* f = std::fopen(path, "w");
  if (!f) return std::string("cannot create ") + path;
  for (size_t k = 0; k < d.vkey.size(); ++k)
    std::fprintf(f, "VERTEX_SE2 %llu %.17g %.17g %.17g\n", (unsigned long long)d.vkey[k], d.vxyt[3 * k], d.vxyt[3 * k + 1],
                 d.vxyt[3 * k + 2]);
  for (size_t k = 0; k < d.pkey.size(); ++k) {
    const double* z = &d.pxyt[3 * k];
    const double* o = &d.pinfo[6 * k];
    std::fprintf(f, "EDGE_PRIOR_SE2 %llu %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", (unsigned long long)d.pkey[k],
                 z[0], z[1], z[2], o[0], o[1], o[2], o[3], o[4], o[5]);
  }
  for (size_t k = 0; k < d.e1.size(); ++k) {
    const double* z = &d.ez[3 * k];
    const double* o = &d.einfo[6 * k];
    std::fprintf(f, "EDGE_SE2 %llu %llu %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", (unsigned long long)d.e1[k],
                 (unsigned long long)d.e2[k], z[0], z[1], z[2], o[0], o[1], o[2], o[3], o[4], o[5]);
  }
  const bool ok = std::fflush(f) == 0;
  std::fclose(f);
  return ok ? std::string() : std::string("write error on ") + path;
}

}  // namespace gs
