/* libgraphslam_cuda — B200-native (sm_100a) pose-graph optimisation backend, C ABI.
 *
 * Drop-in boundary for the optimiser behind Sergimech/GraphSLAM's `src/graph` node.  The node keeps
 * its verbs (prior_factor / new_factor / loop_factor / solve, reference
 * src/graph/src/graph.cpp:27-132); what it hands to GTSAM today
 * (gtsam::NonlinearFactorGraph + gtsam::Values + LevenbergMarquardtOptimizer, graph.cpp:6-7,119)
 * it hands to this library instead.  Each entry point cites the reference call it replaces.
 *
 * Conventions
 *  - plain pointers and sizes only; the caller owns every pointer it passes, inputs are copied
 *    before return; no exceptions cross the ABI; no global state (graph.cpp:5-10's globals live in
 *    gs_graph).
 *  - poses / deltas are (x, y, theta) doubles; covariances are float64[9] row-major exactly as
 *    common/Pose2DWithCovariance.msg:2 and graph.hpp:45-58 (covariance_to_eigen).
 *  - keys are arbitrary uint64 (gtsam::Key); dense index = rank in sorted key order
 *    (= key-1 for the reference's ids, graph.cpp:31,48,68,75).
 *  - a gs_graph is single-writer (the reference runs on one ros::spin() thread, graph.cpp:214);
 *    all device work runs on one per-graph CUDA stream; gs_optimize is synchronous on return.
 *  - there is NO CPU fallback: without a CUDA device gs_create returns GS_ECUDA.
 */
#ifndef GRAPHSLAM_CUDA_H
#define GRAPHSLAM_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_VERSION_MAJOR 0
#define GS_VERSION_MINOR 1

typedef struct gs_graph gs_graph; /* opaque; owns all host and device memory */

typedef enum {
  GS_OK = 0,
  GS_EINVAL = 1,        /* bad argument */
  GS_ENOTSPD = 2,       /* covariance not symmetric positive definite (GTSAM would throw, graph.cpp:45,83,103) */
  GS_EDUPKEY = 3,       /* gs_add_vertex with an existing key (gtsam::Values::insert would throw) */
  GS_ENOKEY = 4,        /* factor or query names an unknown key */
  GS_ECUDA = 5,         /* CUDA runtime failure, see gs_last_error */
  GS_ENCCL = 6,         /* NCCL failure (multi-GPU) */
  GS_ENOTCONVERGED = 7, /* PCG hit pcg_max_iters before pcg_rel_tol */
  GS_ENAN = 8,          /* non-finite chi2 */
  GS_ENOMEM = 9         /* host allocation failed (std::bad_alloc caught at the boundary; no exception crosses the ABI) */
} gs_status;

enum { GS_JACOBIAN_GTSAM_DEFAULT = 0, GS_JACOBIAN_EXACT = 1 };

/* gs_params.solver_flags: the few real choices of execution path (all give the same bits; the defaults are the
 * measured winners).  Fixed at gs_create; nothing in the library reads the environment. */
enum {
  GS_SOLVER_NO_CUDA_GRAPH = 1 << 0,  /* launch the PCG kernels one by one instead of replaying the captured batch */
  GS_SOLVER_NCCL_PER_OP = 1 << 1,    /* multi-GPU: ncclSend/Recv + all-gather per exchange instead of stores into peer memory */
  GS_SOLVER_NO_PERSISTENT = 1 << 2,  /* small graphs: per-kernel iteration instead of the cooperative persistent kernel */
  GS_SOLVER_DIR_KERNEL = 1 << 3,     /* multi-GPU: keep the direction update p = z + beta p as a kernel of its own (three kernels
                                        per iteration); default: formed on the fly inside the product above 200k rows per rank */
  GS_SOLVER_DIR_ON_THE_FLY = 1 << 4, /* multi-GPU: force the on-the-fly form (two kernels per iteration) */
  GS_SOLVER_BLOCK_JACOBI_ONLY = 1 << 5 /* PCG with the 3x3 block-Jacobi preconditioner alone.  Default (lambda*I damping,
                                        >= 1024 poses): block-Jacobi plus a coarse-space correction over aggregates of poses
                                        (rigid-body modes), which cuts the PCG iterations of large graphs ~20x.  The solution
                                        of the linear system is the same to pcg_rel_tol either way. */
};

/* Defaults == gtsam::LevenbergMarquardtParams defaults, i.e. what graph.cpp:119 runs with. */
typedef struct gs_params {
  int max_iterations;        /* 100   */
  double rel_error_tol;      /* 1e-5  */
  double abs_error_tol;      /* 1e-5  */
  double error_tol;          /* 0     */
  double lambda_initial;     /* 1e-5  */
  double lambda_factor;      /* 10    */
  double lambda_upper;       /* 1e5   */
  double lambda_lower;       /* 0     */
  double min_model_fidelity; /* 1e-3  */
  int diagonal_damping;      /* 0 (lambda*I); 1 = lambda*diag(H) clipped to [1e-6,1e32] */
  int jacobian_mode;         /* GS_JACOBIAN_GTSAM_DEFAULT */
  double pcg_rel_tol;        /* 1e-12: |r|_2 <= tol*|g|_2 */
  int pcg_max_iters;         /* 100000 */
  int device;                /* CUDA device ordinal, -1 = current */
  int rank;                  /* multi-GPU: this process's rank (0)      */
  int world_size;            /* multi-GPU: number of ranks (1)          */
  int solver_flags;          /* GS_SOLVER_* bits (0)                    */
  int coarse_aggregate_size; /* two-level preconditioner: poses per aggregate; 0 = automatic (about sqrt(N), >= 48) */
  int peer_wait_limit_ms;    /* multi-GPU, peer-memory path: how long a kernel waits for another GPU before the solve fails
                                with GS_ENCCL (re-run gs_optimize to recover); 0 = 30000 */
  int reserved[5];
} gs_params;

typedef struct gs_stats {
  int iterations;            /* accepted outer LM iterations                                  */
  int outer_loops;           /* linearisations performed (= iterate() calls)                  */
  int inner_iterations;      /* tryLambda calls (linear solves)                               */
  int converged;             /* 1 if a GTSAM stop test fired, 0 if max_iterations             */
  double chi2_initial;       /* 2 * graph.error(initial)                                      */
  double chi2_final;
  double lambda_final;
  int64_t pcg_iterations_total;
  double ms_structure;       /* structure build + upload (0 if topology unchanged)            */
  double ms_linearize;       /* linearise + assemble kernels, summed                          */
  double ms_solve;           /* PCG, summed                                                   */
  double ms_update;          /* retract + chi2 + model-fidelity kernels, summed               */
  double ms_total;           /* wall time of gs_optimize                                      */
  double edges_per_sec;      /* E * outer_loops / ms_linearize                                */
  double gn_iters_per_sec;   /* outer_loops / (ms_total - ms_structure)                       */
  double gpu_launches;       /* kernels launched that did work (no-op graph tail excluded)    */
  double ms_device;          /* CUDA-event time on the graph's stream, first to last kernel   */
  double h2d_bytes;          /* host->device bytes copied by this call                        */
  double d2h_bytes;          /* device->host bytes copied by this call                        */
  double p2p_wait_ticks[3];  /* peer-memory path: SM clock ticks CTA 0 spent waiting for other GPUs in the product /  */
  double p2p_waits;          /*   update / direction kernels during this call, and the iterations they cover        */
  double ms_structure_parts[4]; /* ms_structure split: host structure build, staging + upload, communicator setup
                                   (multi-GPU: peer-memory windows, first NCCL exchanges), CUDA-graph capture        */
  double ms_coarse_setup;    /* two-level preconditioner: coarse-matrix products, Cholesky and inverse (host wall time,
                                part of ms_solve); does not shrink with the number of GPUs                            */
} gs_stats;

void gs_default_params(gs_params* p);
int gs_version(void);

/* lifecycle — replaces the file-scope `graph` / `initial` objects, graph.cpp:6-7 */
gs_status gs_create(gs_graph** out, const gs_params* params /* NULL = defaults */);
/* multi-GPU, one process per GPU: every rank makes the SAME sequence of gs_add_* / gs_optimize / gs_get_poses
 * calls with the same data; the library keeps the block rows of this rank's vertex range on params->device and
 * exchanges halo poses / PCG direction vectors and dot-product segment sums over NCCL (libnccl.so.2 resolved at
 * run time).  `nccl_id` is the ncclUniqueId obtained once by gs_nccl_unique_id on one rank and handed to all
 * (any transport: MPI, torch.distributed, a file).  world_size 1, 2, 4 or 8; results are bitwise identical for
 * every world size.  gs_get_poses / gs_get_pose / gs_optimize / gs_chi2 are collective calls. */
gs_status gs_nccl_unique_id(void* out, size_t cap, size_t* n /* bytes needed/written */);
gs_status gs_create_distributed(gs_graph** out, const gs_params* params /* rank, world_size, device set */,
                                const void* nccl_id, size_t id_bytes);
gs_status gs_destroy(gs_graph* g);
gs_status gs_set_params(gs_graph* g, const gs_params* params);

/* initial.insert(key, Pose2)                                   graph.cpp:58, 86 */
gs_status gs_add_vertex(gs_graph* g, uint64_t key, const double pose_xyt[3]);
/* graph.add(PriorFactor<Pose2>(key, pose, Gaussian::Covariance(Q)))   graph.cpp:45, 57 */
gs_status gs_add_prior(gs_graph* g, uint64_t key, const double pose_xyt[3], const double cov9[9]);
/* graph.add(BetweenFactor<Pose2>(k1, k2, Pose2(dx,dy,dth), Gaussian::Covariance(Q)))
 *                                                               graph.cpp:83-92 (odometry), 103-111 (loop) */
gs_status gs_add_between(gs_graph* g, uint64_t k1, uint64_t k2, const double delta_xyt[3], const double cov9[9]);
/* bulk form of the two calls above (nv vertices then ne between-factors), same semantics */
gs_status gs_add_batch(gs_graph* g, size_t nv, const uint64_t* keys, const double* poses_xyt, size_t ne,
                       const uint64_t* k1, const uint64_t* k2, const double* deltas_xyt, const double* cov9s);

/* the same two factors with noiseModel::Gaussian::Information(I) instead of Covariance(Q);
 * info6 = upper triangle I11 I12 I13 I22 I23 I33 (what a g2o EDGE_SE2 line carries) */
gs_status gs_add_prior_information(gs_graph* g, uint64_t key, const double pose_xyt[3], const double info6[6]);
gs_status gs_add_between_information(gs_graph* g, uint64_t k1, uint64_t k2, const double delta_xyt[3],
                                     const double info6[6]);

/* g2o / TORO 2-D text format (VERTEX_SE2 / EDGE_SE2 / EDGE_PRIOR_SE2; TORO VERTEX2 / EDGE2).  The
 * reference has no persistence: its graph lives in process globals (graph.cpp:5-10) and is lost on
 * exit.  gs_load_g2o appends the file's vertices, priors and edges (all-or-nothing); gs_save_g2o
 * writes the whole graph with the current values at full precision (%.17g), so save -> load
 * reproduces the graph bit for bit (checkpoint / resume). */
typedef struct gs_g2o_counts {
  size_t vertices, edges, priors;
  size_t fixed;         /* ids on FIX lines: reported, not applied (anchor with gs_add_prior as graph.cpp:57 does) */
  size_t skipped_lines; /* lines with a tag this loader does not know */
} gs_g2o_counts;
gs_status gs_load_g2o(gs_graph* g, const char* path, gs_g2o_counts* counts /* nullable */);
gs_status gs_save_g2o(gs_graph* g, const char* path);

/* upload pending vertices/factors and (re)build the sparse structure now; gs_optimize does this
 * lazily.  Separates one-off topology work from the per-iteration path for measurement. */
gs_status gs_prepare(gs_graph* g);
/* wall time of the last structure build in ms: [0] host structure (K6), [1] staging + upload, [2] communicator setup
 * (multi-GPU), [3] CUDA-graph capture */
gs_status gs_get_structure_timing(const gs_graph* g, double ms[4]);

/* gtsam::LevenbergMarquardtOptimizer(graph, initial).optimize(); result kept as the next initial
 * values (warm start)                                           graph.cpp:119, 130 */
gs_status gs_optimize(gs_graph* g, gs_stats* stats /* nullable */);

/* poses_opti.at<Pose2>(id).x()/y()/theta()                      graph.cpp:123-125 */
gs_status gs_get_pose(gs_graph* g, uint64_t key, double out_xyt[3]);
/* all poses in sorted-key order, AoS (x,y,theta), theta in (-pi,pi]; keys/xyt may be NULL */
gs_status gs_get_poses(gs_graph* g, size_t cap, uint64_t* keys, double* xyt, size_t* n);
/* overwrite current values (initial.update)                     */
gs_status gs_set_poses(gs_graph* g, size_t n, const uint64_t* keys, const double* xyt);

/* Rank-local transfers (several GPUs: what scales; one GPU: the same as the calls above without the key lookups).
 * The library keeps the poses in sorted-key ("dense") order and a rank owns a contiguous range of them.  With
 * gs_get_poses every rank receives ALL poses (a broadcast per rank, 32 N bytes over every rank's PCIe link, N
 * conversions on every rank's host); with these a rank moves its own rows only — 32 N / P bytes each way — and the
 * neighbours' rows it needs as ghosts are refreshed device to device.  Valid once the structure is built (gs_prepare /
 * gs_optimize after the last gs_add_*).  No counterpart in graph.cpp (one process, one optimiser); the write-back loop
 * graph.cpp:122-125 maps to gs_get_local_poses on each rank. */
/* the dense rows [first, first + count) this rank owns */
gs_status gs_local_range(gs_graph* g, size_t* first, size_t* count);
/* keys and (x, y, theta) of the owned rows, in dense order; keys / xyt may be NULL; NOT collective */
gs_status gs_get_local_poses(gs_graph* g, size_t cap, uint64_t* keys, double* xyt, size_t* count);
/* new values for the owned rows (n == count, dense order).  COLLECTIVE on several GPUs: every rank sets its rows, then
 * the ghost rows are exchanged.  Non-finite input: GS_EINVAL, the rank's rows stay as they were. */
gs_status gs_set_local_poses(gs_graph* g, size_t n, const double* xyt);
/* 2 * graph.error(values) at the current values                 */
gs_status gs_chi2(gs_graph* g, double* chi2);
gs_status gs_num_vertices(const gs_graph* g, size_t* n);
/* cumulative bytes the library has copied host->device / device->host for this graph (measurement) */
gs_status gs_get_transfer_bytes(const gs_graph* g, double* h2d, double* d2h);
gs_status gs_num_factors(const gs_graph* g, size_t* n);          /* graph.nrFactors(), graph.cpp:60 */

/* ---- parity hooks (integer structure must match the oracle bit-exactly) ---- */
/* sorted unique 3x3 block pattern of H incl. the diagonal; pass NULLs to query nnzb */
gs_status gs_get_pattern(gs_graph* g, int32_t* rowptr /* N+1 */, int32_t* colidx /* nnzb */, size_t* nnzb);
/* linearise at the current values: blocks[nnzb*9] row-major on the pattern above, b[3N] = sum J^T Omega e */
gs_status gs_get_system(gs_graph* g, double* blocks, double* b, double* chi2);
/* solve (H + lambda*I) delta = -b at the current linearisation by block-Jacobi PCG; delta[3N] */
gs_status gs_solve_linear(gs_graph* g, double lambda, double* delta, int* pcg_iters, double* rel_residual);
/* gtsam::Marginals(graph, poses_opti).marginalCovariance(id)     graph.cpp:120, 126-127 (commented out
 * in the reference): the 3x3 covariance of one pose in its own tangent space = the (id, id) block of
 * (J^T Omega J)^-1 at the current values.  Three undamped PCG solves per pose on the device; cov9 row-major.
 * On a distributed graph every rank calls it with the same keys and receives the same values. */
gs_status gs_marginal_covariance(gs_graph* g, uint64_t key, double cov9[9]);
gs_status gs_marginal_covariances(gs_graph* g, size_t n, const uint64_t* keys, double* cov9s);

/* per-outer-iteration trace of the last gs_optimize: errors[n+1] (= chi2/2, [0] initial), lambdas[n], tries[n] */
gs_status gs_get_trace(gs_graph* g, size_t cap, double* errors, double* lambdas, int32_t* tries, size_t* n);

/* ---- measurement hook: run `iters` real PCG iterations on the current linearisation (un-graphed, a
 * CUDA-event pair around every launch on the graph's own stream) and return the average duration of
 * each kernel in ms: ms[0] = BSR SpMV, ms[1] = fused x/r/z update + dots, ms[2] = direction update.
 * bytes[0..2] receive the algorithmic bytes of one launch (SURVEY.md section 8(d) formulas). */
gs_status gs_profile_pcg(gs_graph* g, double lambda, int iters, double ms[3], double bytes[3]);
/* same for the once-per-outer-iteration kernels: ms[0] = linearise+assemble, ms[1] = retract, ms[2] = chi2 */
gs_status gs_profile_outer(gs_graph* g, int reps, double ms[3], double bytes[3]);

const char* gs_last_error(const gs_graph* g);

/* host helpers: graph.hpp:30-43 restated bit-for-bit (input_1's translation dropped, one-sided fmod wrap) */
void gs_compose_reference(const double a_xyt[3], const double b_xyt[3], double out_xyt[3]);
/* the correct SE(2) a (+) b (gtsam::Pose2::compose), theta wrapped by atan2 */
void gs_compose(const double a_xyt[3], const double b_xyt[3], double out_xyt[3]);

/* closest_keyframe service, graph.cpp:146-178: nearest (Euclidean x,y) vertex to `key` among all but
 * the last `skip` vertices (insertion order), ties -> lowest index (strict '<', graph.cpp:162).
 * Returns GS_ENOKEY if fewer than skip+1 vertices exist (service returns false). */
gs_status gs_closest_keyframe(gs_graph* g, uint64_t key, int skip, uint64_t* out_key, double* out_dist);

/* The same query for callers that must not allocate (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * It calls no cudaMalloc / cudaFree / allocating library routine: the reduction's scratch (per-CTA partial results,
 * result, last-CTA counter) lives in a fixed file-scope device array that every GPU loading the library holds until
 * process exit (64 KB), which bounds the candidate count: at most GS_CLOSEST_NOALLOC_MAX_POSES keyframes
 * (GS_EINVAL with a message beyond that).  Requirements, each answered with GS_EINVAL and a message instead of an
 * allocation: a single-rank graph that is prepared with its values on the device (gs_prepare / gs_optimize after the
 * last gs_add_* or gs_set_poses); keys inserted in increasing order, or one earlier gs_closest_keyframe call since the
 * last topology change.  Same result as gs_closest_keyframe, bit for bit.  All calls (of all graphs on one GPU) share
 * the array: calls that overlap on the GPU race; calls on one graph, or serialised by the caller, do not. */
#define GS_CLOSEST_NOALLOC_MAX_POSES 8388608
gs_status gs_closest_keyframe_noalloc(gs_graph* g, uint64_t key, int skip, uint64_t* out_key, double* out_dist);

#ifdef __cplusplus
}
#endif
#endif /* GRAPHSLAM_CUDA_H */
