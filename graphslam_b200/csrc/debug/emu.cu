// TEST-ONLY emulator of the multi-rank device path (include/graphslam_cuda_debug.h: gs_emu_run).
//
// It runs, on the CPU and over V virtual ranks, the same structure build (structure.hpp), the same
// per-row bodies (rows.cuh) and the same tile -> segment -> total reduction tree as the CUDA
// kernels, with the NCCL halo exchange and all-gathers replaced by plain copies between the
// virtual ranks' arrays.  The CPU test-suite uses it to check the partition / halo / indexing
// logic against the oracle and to check that the result is bitwise independent of the number of
// ranks.  It is NOT a fallback: no product entry point (gs_create, gs_optimize, ...) reaches it,
// and those fail with GS_ECUDA when no CUDA device is present.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../../include/graphslam_cuda_debug.h"
#include "../rows.cuh"

namespace {

using gs::LocalStructure;

struct VRank {
  LocalStructure S;
  int n_own = 0, n_loc = 0;
  std::vector<int2> refs;
  std::vector<int4> cw;
  gs::FactorStage fs;
  std::vector<double4> pr_pose, pose, pose_try;
  std::vector<double> pr_om, Ho, Hd, Minv, b, x, r, z, p, Ap, part;
  size_t hstride = 1;
  gs::AsmView av(const double4* ps) {
    gs::AsmView a;
    a.cs.base = S.csell.base.data();
    a.cs.width = S.csell.width.data();
    a.c = cw.data();
    a.c_first = S.c_first.data();
    a.f.ij = fs.ij.data();
    a.f.data = fs.data.data();
    a.f.fstride = fs.fstride;
    a.pr.n = S.P_loc;
    a.pr.row = S.pr_row.data();
    a.pr.pose = pr_pose.data();
    a.pr.om = pr_om.data();
    a.pose = ps;
    a.Ho = Ho.data();
    a.hstride = hstride;
    a.Hd = Hd.data();
    a.b = b.data();
    a.n_own = n_own;
    return a;
  }
  gs::SpmvView sv() {
    gs::SpmvView a;
    a.rs.base = S.rsell.base.data();
    a.rs.width = S.rsell.width.data();
    a.refs = refs.data();
    a.Ho = Ho.data();
    a.hstride = hstride;
    a.Hd = Hd.data();
    a.dscale = nullptr;
    a.n_own = n_own;
    return a;
  }
  gs::PcgView pv() {
    gs::PcgView a;
    a.b = b.data();
    a.Minv = Minv.data();
    a.x = x.data();
    a.r = r.data();
    a.z = z.data();
    a.p = p.data();
    a.Ap = Ap.data();
    a.n_own = n_own;
    return a;
  }
  int global_of(int l) const { return l < n_own ? S.v0 + l : S.ghosts[l - n_own]; }
};

struct World {
  std::vector<VRank> R;
  int V = 1;
  // halo exchange: ghost slots of every rank <- owner's values (what the grouped ncclSend/ncclRecv do),
  // going through the same send lists / ghost ranges the device path uses
  template <typename GetVec>
  void halo(GetVec vec, int width) {
    for (VRank& me : R)
      for (const gs::HaloPeer& hp : me.S.peers) {
        VRank& peer = R[hp.rank];
        // find the peer's matching entry (its send list to me)
        for (const gs::HaloPeer& back : peer.S.peers) {
          if (back.rank != me.S.rank) continue;
          double* dst = vec(me) + ((size_t)me.n_own + hp.ghost_off) * width;
          const double* src = vec(peer);
          for (int k = 0; k < back.send_cnt; ++k)
            for (int c = 0; c < width; ++c) dst[(size_t)k * width + c] = src[(size_t)peer.S.send_rows[back.send_off + k] * width + c];
        }
      }
  }
  // per-row values -> tile partials -> segment sums -> total, all in fixed order
  template <typename RowFn>
  double reduce(RowFn fn) {
    double segs[gs::kSegments];
    const int spr = gs::kSegments / V;
    for (VRank& me : R) {
      const int nt = std::max(me.S.ntiles, 1);
      me.part.assign(nt, 0.0);
      for (int t = 0; t < me.S.ntiles; ++t) {
        double s = 0.0;
        for (int r = t * gs::kTile; r < std::min(me.n_own, (t + 1) * gs::kTile); ++r) s += fn(me, r);
        me.part[t] = s;
      }
      for (int k = 0; k < spr; ++k) {
        double s = 0.0;
        for (int t = me.S.seg_tile[k]; t < me.S.seg_tile[k + 1]; ++t) s += me.part[t];
        segs[me.S.rank * spr + k] = s;
      }
    }
    double tot = 0.0;
    for (int g = 0; g < gs::kSegments; ++g) tot += segs[g];
    return tot;
  }
};

}  // namespace

extern "C" gs_status gs_host_partition(int32_t n, int32_t e, const int32_t* ei, const int32_t* ej, int world, int rank,
                                       int32_t info[8], int32_t* ghosts, int32_t* peers, int32_t* send_rows) {
  if (n <= 0 || e < 0 || !info || rank < 0 || rank >= world) return GS_EINVAL;
  if (!(world == 1 || world == 2 || world == 4 || world == 8)) return GS_EINVAL;
  std::vector<int> perm;
  gs::canonical_order(n, e, ei, ej, perm);
  LocalStructure S;
  gs::build_local_structure(n, e, ei, ej, perm, 0, nullptr, world, rank, S);
  info[0] = S.v0;
  info[1] = S.v1;
  info[2] = S.n_ghost;
  info[3] = (int32_t)S.peers.size();
  info[4] = (int32_t)S.send_rows.size();
  info[5] = S.n_pairs;
  info[6] = S.nnzo_own;
  info[7] = S.ntiles;
  if (ghosts) std::memcpy(ghosts, S.ghosts.data(), sizeof(int32_t) * S.ghosts.size());
  if (peers)
    for (size_t k = 0; k < S.peers.size(); ++k) {
      const gs::HaloPeer& hp = S.peers[k];
      const int32_t row[5] = {hp.rank, hp.send_off, hp.send_cnt, hp.ghost_off, hp.ghost_cnt};
      std::memcpy(peers + 5 * k, row, sizeof row);
    }
  if (send_rows)
    for (size_t k = 0; k < S.send_rows.size(); ++k) send_rows[k] = S.v0 + S.send_rows[k];  // global indices
  return GS_OK;
}

static gs_status emu_run(const gs_emu_problem* pb, int world, int jacobian_mode, double lambda, int pcg_max_iters,
                         double pcg_rel_tol, int form, gs_emu_result* out) {
  if (!pb || !out || pb->n <= 0) return GS_EINVAL;
  if (!(world == 1 || world == 2 || world == 4 || world == 8)) return GS_EINVAL;
  const int N = pb->n, E = pb->e, P = pb->np;
  std::vector<int> perm;
  gs::canonical_order(N, E, pb->ei, pb->ej, perm);
  World W;
  W.V = world;
  W.R.resize(world);
  for (int q = 0; q < world; ++q) {
    VRank& me = W.R[q];
    gs::build_local_structure(N, E, pb->ei, pb->ej, perm, P, pb->prior_idx, world, q, me.S);
    const LocalStructure& S = me.S;
    me.n_own = S.n_own;
    me.n_loc = S.n_own + S.n_ghost;
    const int FS = S.csell.slots;
    me.fs.init(FS);
    me.cw.resize(S.c_code.size());
    for (size_t c = 0; c < me.cw.size(); ++c) me.cw[c] = make_int4((int)S.c_code[c], S.c_blk[c], S.c_col[c], 0);
    me.refs.resize(S.r_col.size());
    for (size_t c = 0; c < me.refs.size(); ++c) me.refs[c] = make_int2(S.r_col[c], S.r_blk[c]);
    for (int s = 0; s < FS; ++s) {
      const int k = S.f_src[s];
      if (k < 0) continue;
      me.fs.set(s, S.f_i[s] | (S.f_count[s] ? 0 : gs::kNoCountBit), S.f_j[s], pb->z + 3 * (size_t)k, pb->om + 6 * (size_t)k);
    }
    me.pr_pose.resize(std::max(S.P_loc, 1));
    me.pr_om.assign(6 * (size_t)std::max(S.P_loc, 1), 0.0);
    for (int k = 0; k < S.P_loc; ++k) {
      const int s = S.pr_src[k];
      me.pr_pose[k] = make_double4(pb->prior_xyt[3 * s], pb->prior_xyt[3 * s + 1], std::cos(pb->prior_xyt[3 * s + 2]),
                                   std::sin(pb->prior_xyt[3 * s + 2]));
      for (int m = 0; m < 6; ++m) me.pr_om[6 * (size_t)k + m] = pb->prior_om[6 * (size_t)s + m];
    }
    me.pose.resize(std::max(me.n_loc, 1));
    me.pose_try.resize(std::max(me.n_loc, 1));
    for (int l = 0; l < me.n_loc; ++l) {
      const int d = me.global_of(l);
      me.pose[l] = make_double4(pb->xyt[3 * d], pb->xyt[3 * d + 1], std::cos(pb->xyt[3 * d + 2]), std::sin(pb->xyt[3 * d + 2]));
    }
    me.hstride = (size_t)std::max(S.rsell.slots, gs::kSlice);
    me.Ho.assign(9 * me.hstride, 0.0);
    const size_t n3 = 3 * (size_t)std::max(me.n_own, 1), nl3 = 3 * (size_t)std::max(me.n_loc, 1);
    me.Hd.assign(2 * n3, 0.0);
    me.Minv.assign(2 * n3, 0.0);
    me.b.assign(n3, 0.0);
    me.r.assign(n3, 0.0);
    me.z.assign(nl3, 0.0);   // ghost room: the two-kernel form exchanges z
    me.Ap.assign(n3, 0.0);
    me.x.assign(nl3, 0.0);
    me.p.assign(nl3, 0.0);
  }

  // ---- K1+K2
  if (form & 2)
    for (VRank& me : W.R)
      if (me.S.n_merged != 0) return GS_EINVAL;   // a pair with two factors needs the run accumulator
  const double chi2 = W.reduce([&](VRank& me, int r) {
    const gs::AsmView a = me.av(me.pose.data());
    if (form & 2)   // the form without run accumulator (k_assemble_nodup); only valid when no pair has two factors
      return jacobian_mode ? gs::assemble_row<1, true>(a, r) : gs::assemble_row<0, true>(a, r);
    return jacobian_mode ? gs::assemble_row<1>(a, r) : gs::assemble_row<0>(a, r);
  });
  out->chi2 = chi2;
  out->chi2_check = W.reduce([&](VRank& me, int r) { return gs::chi2_row<1>(me.av(me.pose.data()), r); });

  // ---- system on the global pattern
  if (out->blocks || out->b) {
    std::vector<int32_t> rp, ci;
    gs::global_pattern(N, E, pb->ei, pb->ej, rp, ci);
    for (VRank& me : W.R) {
      const LocalStructure& S = me.S;
      for (int r = 0; r < me.n_own; ++r) {
        const int g = S.v0 + r;
        if (out->b)
          for (int k = 0; k < 3; ++k) out->b[3 * (size_t)g + k] = me.b[3 * (size_t)r + k];
        if (!out->blocks) continue;
        size_t w = (size_t)rp[g];
        bool diag_done = false;
        auto put_diag = [&]() {
          const size_t n = (size_t)me.n_own;
          const double d[6] = {me.Hd[r], me.Hd[n + r], me.Hd[2 * n + r], me.Hd[3 * n + r], me.Hd[4 * n + r], me.Hd[5 * n + r]};
          const double full[9] = {d[0], d[1], d[2], d[1], d[3], d[4], d[2], d[4], d[5]};
          std::memcpy(out->blocks + 9 * w, full, sizeof full);
          ++w;
          diag_done = true;
        };
        const int wd = S.rsell.width[r / gs::kSlice];
        for (int s = 0; s < wd; ++s) {
          const int q = S.rsell.pos(r, s);
          const int lc = S.r_col[q];
          if (lc < 0) break;
          const int gc = me.global_of(lc);
          if (!diag_done && gc > g) put_diag();
          if (ci[w] != gc) return GS_EINVAL;  // reference list disagrees with the global pattern
          const int word = S.r_blk[q], pos = word >> 1;
          const bool tr = word & 1;
          for (int a = 0; a < 3; ++a)
            for (int c = 0; c < 3; ++c) out->blocks[9 * w + 3 * a + c] = gs::h_entry(me.Ho.data(), me.hstride, pos, a, c, tr);
          ++w;
        }
        if (!diag_done) put_diag();
        if (w != (size_t)rp[g + 1]) return GS_EINVAL;
      }
    }
  }

  // ---- block-Jacobi PCG on (H + lambda I) x = -b
  for (VRank& me : W.R)
    for (int r = 0; r < me.n_own; ++r) gs::minv_row(me.Hd.data(), nullptr, me.Minv.data(), me.n_own, r, lambda);
  double rr = 0.0;
  double rz = W.reduce([&](VRank& me, int r) {
    double a, b;
    gs::pcg_init_row(me.pv(), r, a, b);
    return a;
  });
  rr = W.reduce([&](VRank& me, int r) { return gs::dot_row(me.r.data(), me.r.data(), r); });
  const double rr0 = rr;
  int it = 0, done = (rr == 0.0) ? 1 : 0;
  form &= 1;
  if (form == 1) {
    // the two-kernel form of the peer-memory path and of the persistent small-graph kernel: the direction is not
    // stored before the product; p(c) = z(c) + beta p_prev(c) is formed on the fly (ghost columns from the exchanged
    // z and the locally kept ghost p), the new p goes to a second buffer
    std::vector<std::vector<double>> alt(W.R.size());
    for (size_t q = 0; q < W.R.size(); ++q) alt[q].assign(W.R[q].p.size(), 0.0);
    std::vector<double*> prev(W.R.size()), next(W.R.size());
    for (size_t q = 0; q < W.R.size(); ++q) {
      prev[q] = W.R[q].p.data();
      next[q] = alt[q].data();
      std::fill(W.R[q].p.begin() + 3 * (size_t)W.R[q].n_own, W.R[q].p.end(), 0.0);
    }
    double beta = 0.0;
    while (!done && it < pcg_max_iters) {
      W.halo([](VRank& v) { return v.z.data(); }, 3);
      for (size_t q = 0; q < W.R.size(); ++q) {
        VRank& me = W.R[q];
        const size_t own3 = 3 * (size_t)me.n_own;
        for (size_t e = own3; e < me.p.size(); ++e) next[q][e] = gs::dir_value(me.z[e], beta, prev[q][e]);
      }
      const double pAp = W.reduce([&](VRank& me, int r) {
        const size_t q = (size_t)me.S.rank, own3 = 3 * (size_t)me.n_own;
        return gs::spmv_row_t<true>(me.sv(), prev[q], prev[q] + own3, me.z.data(), me.z.data() + own3, beta, next[q],
                                    me.Ap.data(), r, lambda);
      });
      for (size_t q = 0; q < W.R.size(); ++q) std::swap(prev[q], next[q]);
      if (!(pAp > 0.0)) {
        done = 2;
        break;
      }
      const double alpha = rz / pAp;
      const double nrz = W.reduce([&](VRank& me, int r) {
        double a, b;
        gs::PcgView v = me.pv();
        v.p = prev[(size_t)me.S.rank];
        gs::pcg_update_row(v, r, alpha, a, b);
        return a;
      });
      rr = W.reduce([&](VRank& me, int r) { return gs::dot_row(me.r.data(), me.r.data(), r); });
      beta = nrz / rz;
      rz = nrz;
      ++it;
      if (rr <= pcg_rel_tol * pcg_rel_tol * rr0) done = 1;
      else if (!(rr == rr)) done = 2;
    }
  }
  while (form == 0 && !done && it < pcg_max_iters) {
    W.halo([](VRank& v) { return v.p.data(); }, 3);
    const double pAp = W.reduce([&](VRank& me, int r) { return gs::spmv_row(me.sv(), me.p.data(), me.p.data() + 3 * (size_t)me.n_own, me.Ap.data(), r, lambda); });
    if (!(pAp > 0.0)) {
      done = 2;
      break;
    }
    const double alpha = rz / pAp;
    // r.z and r.r are two reductions over the same rows: run the update once, recompute r.r from r
    const double nrz = W.reduce([&](VRank& me, int r) {
      double a, b;
      gs::pcg_update_row(me.pv(), r, alpha, a, b);
      return a;
    });
    rr = W.reduce([&](VRank& me, int r) { return gs::dot_row(me.r.data(), me.r.data(), r); });
    const double beta = nrz / rz;
    rz = nrz;
    ++it;
    if (rr <= pcg_rel_tol * pcg_rel_tol * rr0) done = 1;
    else if (!(rr == rr)) done = 2;
    if (done) break;
    for (VRank& me : W.R)
      for (int r = 0; r < me.n_own; ++r) gs::pcg_dir_row(me.pv(), r, beta);
  }
  out->pcg_iters = it;
  out->pcg_done = done;
  out->rel_residual = rr0 > 0.0 ? std::sqrt(rr / rr0) : 0.0;

  // ---- model terms, retract, new chi2
  W.halo([](VRank& v) { return v.x.data(); }, 3);
  out->dHd = W.reduce([&](VRank& me, int r) { return gs::spmv_row(me.sv(), me.x.data(), me.x.data() + 3 * (size_t)me.n_own, me.Ap.data(), r, 0.0); });
  out->db = W.reduce([&](VRank& me, int r) { return gs::dot_row(me.x.data(), me.b.data(), r); });
  for (VRank& me : W.R)
    for (int r = 0; r < me.n_own; ++r)
      me.pose_try[r] = gs::retract(me.pose[r], me.x[3 * (size_t)r], me.x[3 * (size_t)r + 1], me.x[3 * (size_t)r + 2]);
  W.halo([](VRank& v) { return reinterpret_cast<double*>(v.pose_try.data()); }, 4);
  out->chi2_new = W.reduce([&](VRank& me, int r) { return gs::chi2_row<2>(me.av(me.pose_try.data()), r); });
  for (VRank& me : W.R)
    for (int r = 0; r < me.n_own; ++r) {
      const int g = me.S.v0 + r;
      if (out->delta)
        for (int k = 0; k < 3; ++k) out->delta[3 * (size_t)g + k] = me.x[3 * (size_t)r + k];
      if (out->xyt_new) {
        out->xyt_new[3 * (size_t)g] = me.pose_try[r].x;
        out->xyt_new[3 * (size_t)g + 1] = me.pose_try[r].y;
        out->xyt_new[3 * (size_t)g + 2] = std::atan2(me.pose_try[r].w, me.pose_try[r].z);
      }
    }
  // structure summary of rank 0 .. world-1 (sums)
  out->n_ghost_total = 0;
  out->n_pairs_total = 0;
  out->halo_send_total = 0;
  out->n_factor_copies = 0;
  out->n_block_copies = 0;
  for (VRank& me : W.R) {
    out->n_ghost_total += me.S.n_ghost;
    out->n_pairs_total += me.S.n_pairs;
    out->halo_send_total += (int)me.S.send_rows.size();
    out->n_factor_copies += me.S.n_factor_copies;
    out->n_block_copies += me.S.n_block_copies;
  }
  return GS_OK;
}

extern "C" gs_status gs_emu_run(const gs_emu_problem* pb, int world, int jacobian_mode, double lambda, int pcg_max_iters,
                                double pcg_rel_tol, gs_emu_result* out) {
  return emu_run(pb, world, jacobian_mode, lambda, pcg_max_iters, pcg_rel_tol, 0, out);
}

// form 1: the PCG iteration with the direction formed on the fly (k_spmvdir_p2p / k_pcg_fused); same bits as form 0
extern "C" gs_status gs_emu_run_form(const gs_emu_problem* pb, int world, int jacobian_mode, double lambda, int pcg_max_iters,
                                     double pcg_rel_tol, int form, gs_emu_result* out) {
  if (form < 0 || form > 3) return GS_EINVAL;   // bit 0: PCG form, bit 1: assembly without run accumulator
  return emu_run(pb, world, jacobian_mode, lambda, pcg_max_iters, pcg_rel_tol, form, out);
}
