/* libgraphslam_cuda — TEST-ONLY entry points (not part of the drop-in boundary).
 *
 * These exercise the library's own host code on a box without a GPU: the structure build, the
 * vertex-range partition and halo lists, and — through a CPU emulator that calls the very same
 * __host__ __device__ per-row functions the CUDA kernels call — one linearise/assemble, a block-Jacobi
 * PCG solve, retract and chi2 over 1, 2, 4 or 8 virtual ranks.  Nothing here is reachable from
 * gs_create / gs_optimize: the product path has no CPU fallback and fails with GS_ECUDA without a
 * device.  Used by tests/ only.
 */
#ifndef GRAPHSLAM_CUDA_DEBUG_H
#define GRAPHSLAM_CUDA_DEBUG_H

#include "graphslam_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* sorted unique block pattern (rowptr[n+1], colidx[nnzb]) and the canonical factor order `perm`
 * of a graph over dense indices 0..n-1; any output may be NULL */
gs_status gs_host_pattern(int32_t n, int32_t e, const int32_t* ei, const int32_t* ej, int32_t* rowptr, int32_t* colidx,
                          size_t* nnzb, int32_t* perm);

/* the aggregates of the two-level preconditioner's coarse space, as gs_prepare builds them (target <= 0: the automatic
 * size max(48, sqrt n)); counts = {aggregates, colours of the distance-2 colouring, target}; agg_of[n], color[aggregates]
 * (size it from a first call) and seg_agg[9] (first aggregate of each fixed reduction segment) may be NULL */
gs_status gs_host_aggregation(int32_t n, int32_t e, const int32_t* ei, const int32_t* ej, int32_t target, int32_t counts[3],
                              int32_t* agg_of, int32_t* color, int32_t* seg_agg);

/* one BetweenFactor evaluated on the host by the inline functions the kernels compile */
gs_status gs_host_between(const double pi[3], const double pj[3], const double z[3], const double cov9[9], int mode,
                          double e[3], double Ai[9], double Aj[9], double Hii[9], double Hij[9], double Hjj[9],
                          double bi[3], double bj[3]);

/* the g2o / TORO reader and writer behind gs_load_g2o / gs_save_g2o, without a graph handle: parse in_path,
 * counts = {vertices, edges, priors, fixed, skipped lines}; array outputs (sized from a first call) and out_path
 * (re-written file) may be NULL */
gs_status gs_host_g2o(const char* in_path, const char* out_path, size_t counts[5], uint64_t* vkey, double* vxyt,
                      uint64_t* e1, uint64_t* e2, double* ez, double* einfo, uint64_t* pkey, double* pxyt, double* pinfo);

/* vertex-range partition of rank `rank` of `world`: info = {v0, v1, n_ghost, n_peers, n_send, n_pairs, nnzo_own,
 * ntiles}; ghosts[n_ghost] (global, sorted); peers[5*n_peers] = {rank, send_off, send_cnt, ghost_off, ghost_cnt};
 * send_rows[n_send] (global indices, grouped by peer).  Array outputs may be NULL (size them with n). */
gs_status gs_host_partition(int32_t n, int32_t e, const int32_t* ei, const int32_t* ej, int world, int rank,
                            int32_t info[8], int32_t* ghosts, int32_t* peers, int32_t* send_rows);

typedef struct gs_emu_problem {
  int32_t n, e, np;            /* vertices, between-factors, priors (dense indices 0..n-1) */
  const int32_t* ei;           /* id_1 per factor */
  const int32_t* ej;           /* id_2 per factor */
  const double* z;             /* 3 per factor */
  const double* om;            /* information upper triangle, 6 per factor: 00 01 02 11 12 22 */
  const int32_t* prior_idx;
  const double* prior_xyt;     /* 3 per prior */
  const double* prior_om;      /* 6 per prior */
  const double* xyt;           /* 3 per vertex: linearisation point */
} gs_emu_problem;

typedef struct gs_emu_result {
  double* blocks;              /* nnzb*9 on gs_host_pattern's pattern, or NULL */
  double* b;                   /* 3n or NULL */
  double* delta;               /* 3n or NULL: PCG solution of (H + lambda I) delta = -b */
  double* xyt_new;             /* 3n or NULL: retract(xyt, delta) */
  double chi2, chi2_check;     /* from the assembly pass / from the chi2 pass */
  double chi2_new;             /* chi2 at xyt_new */
  double dHd, db;              /* delta.H.delta (undamped), delta.b */
  double rel_residual;
  int32_t pcg_iters, pcg_done;
  int32_t n_ghost_total, n_pairs_total, halo_send_total;
  int32_t n_factor_copies, n_block_copies; /* cells that keep a copy because sharing would not coalesce */
} gs_emu_result;

/* one outer-iteration body over `world` virtual ranks (1, 2, 4 or 8) on the CPU */
gs_status gs_emu_run(const gs_emu_problem* pb, int world, int jacobian_mode, double lambda, int pcg_max_iters,
                     double pcg_rel_tol, gs_emu_result* out);

/* the same with the PCG iteration in the form the two-kernel peer-memory path and the persistent small-graph kernel
 * use (form bit 0: direction formed on the fly inside the product, z exchanged instead of p; form bit 1: assembly
 * without the run accumulator, GS_EINVAL when some vertex pair carries two factors); form 0 = gs_emu_run */
gs_status gs_emu_run_form(const gs_emu_problem* pb, int world, int jacobian_mode, double lambda, int pcg_max_iters,
                          double pcg_rel_tol, int form, gs_emu_result* out);

#ifdef __cplusplus
}
#endif
#endif /* GRAPHSLAM_CUDA_DEBUG_H */
