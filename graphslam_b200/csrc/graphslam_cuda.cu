// libgraphslam_cuda: C ABI (include/graphslam_cuda.h) + Levenberg-Marquardt driver.
//
// The driver restates gtsam::LevenbergMarquardtOptimizer::optimize() with default parameters —
// what reference src/graph/src/graph.cpp:119 calls — as: k_assemble -> [k_minv -> PCG] per
// tryLambda -> k_retract + k_chi2 -> accept/reject on three scalars read back per try.
// The PCG loop itself never synchronises with the host: coefficients live in device memory and
// iterations are replayed from a CUDA graph in batches, a `done` flag turning the tail into no-ops.
//
// Multi-GPU (one process per GPU, gs_create_distributed): every rank receives the same graph
// through the same calls and keeps the block rows of its vertex range.  NCCL carries the halo
// exchange of p (per SpMV) and of the poses (per retract), and all-gathers of the segment sums of
// the dot products, which every rank then adds in the same fixed order.
#include <cuda_runtime.h>
#include <cublas_v2.h>
#include <cusolverDn.h>
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <limits>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/graphslam_cuda.h"
#include "comm.hpp"
#include "g2o.hpp"
#include "host_util.hpp"
#include "kernels.cuh"
#include "kernels_coarse.cuh"
#include "kernels_fused.cuh"
#include "kernels_p2p.cuh"

namespace {

using gs::LocalStructure;
using gs::Scalars;
using gs::spd_information;
using gs::spd_upper;

constexpr int kPcgBatch = 32;  // PCG iterations per CUDA-graph launch
constexpr int kCoarseMinPoses = 1024;  // smaller graphs keep block-Jacobi PCG (persistent kernel)
constexpr int kFusedBatch = 128;  // PCG iterations per launch of the persistent small-graph kernel
static_assert(kFusedBatch % 2 == 0, "an even batch leaves the latest direction in the graph's own p vector");

// NVTX range per stage (prepare / linearise / solve / coarse setup / update): visible in nsys / ncu timelines
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

double now_ms() {
  using namespace std::chrono;
  return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  // grows geometrically: a graph that gains a hundred poses per update (incremental mode) must not free and
  // reallocate its forty buffers every time (cudaFree / cudaMalloc synchronise the device)
  cudaError_t alloc(size_t count) {
    if (count <= n && p) return cudaSuccess;
    const bool growing = p != nullptr;
    release();
    if (count == 0) return cudaSuccess;
    n = growing ? count + count / 4 + 64 : count;
    return cudaMalloc(&p, n * sizeof(T));
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  ~DevBuf() { release(); }
};

}  // namespace

struct gs_graph {
  gs_params prm;
  std::string err;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_begin = nullptr, ev_end = nullptr;
  double h2d_bytes = 0.0, d2h_bytes = 0.0;  // cumulative bytes copied by the library

  // ---- multi-GPU
  int world = 1, rank = 0;
  ncclComm_t comm = nullptr;
  // execution-path choices: gs_params.solver_flags (fixed at gs_create), never the environment
  bool use_graph = true;                  // !GS_SOLVER_NO_CUDA_GRAPH
  bool dir_otf = false;                   // direction formed on the fly inside the product: two kernels per iteration
                                          // (k_spmvdir[_p2p] + update) instead of three; decided in prepare()
  int ghost_chunk = 0;                    // ghost doubles of p advanced per halo tile in k_spmvdir_p2p
  bool fused_wanted = true;               // !GS_SOLVER_NO_PERSISTENT
  bool fused = false;                     // this graph's tiles are all co-resident: one persistent kernel per batch
  int fused_capacity = 0;                 // CTAs of k_pcg_fused the device holds at once
  DevBuf<gs::GridBarrier> d_bar;
  DevBuf<double> d_p_alt;                 // second direction buffer of the persistent kernel
  // peer-memory path (world > 1): collectives fused into the PCG kernels over NVLink
  bool p2p_wanted = true;                 // !GS_SOLVER_NCCL_PER_OP
  bool p2p = false;                       // windows exchanged, peer-memory kernels in use
  char* win = nullptr;                    // this rank's exported window: header + p
  long long spin_limit_ticks = 60000000000LL;   // peer_wait_limit_ms x SM clock: how long a kernel waits for another GPU
  size_t win_bytes = 0;
  size_t win_ghost_off = gs::kWindowHeaderBytes;   // where the ghost rows of p start inside a window (same on every rank)
  int win_generation = 0;
  char* peer_win[gs::kMaxRanks] = {nullptr};   // peer mappings (allocation base + offset)
  void* peer_base[gs::kMaxRanks] = {nullptr};  // what cudaIpcOpenMemHandle returned
  int peer_generation[gs::kMaxRanks] = {0};
  unsigned int nbr_mask = 0;
  DevBuf<unsigned int> d_seq;             // [0] sequence base, [32] ticket (a line each)
  int n_push_tiles = 0;                   // tiles that send halo rows
  DevBuf<int4> d_hp_order;                 // CTA -> tile for k_dir_p2p (sending tiles first)
  DevBuf<int> d_tile_halo;                // [tile] 1 when the tile's rows reference ghost columns
  DevBuf<gs::P2PWindow*> d_winptr;        // [rank] window pointers as the kernels see them
  DevBuf<int> d_hp_row;
  DevBuf<double*> d_hp_dst;
  DevBuf<char> d_xchg;
  double* pvec = nullptr;                 // direction vector p (owned rows, then the NCCL path's ghost rows)
  double* pghost = nullptr;               // ghost rows of p: behind pvec, or in the window

  // ---- host copy of the graph, insertion order (what graph.cpp keeps in `graph` and `initial`)
  std::vector<uint64_t> keys;
  std::vector<double> xyt;  // current values, 3 per vertex, insertion order
  std::unordered_map<uint64_t, int> key2ins;
  bool keys_sequential = true;  // keys[k] == keys[0] + k  -> no hash lookups needed
  std::vector<int> e1, e2;      // insertion indices of id_1, id_2
  std::vector<double> ez;       // 3 per factor
  std::vector<double> eom;      // information, 6 per factor
  std::vector<int> pr_v;
  std::vector<double> pr_xyt, pr_om;

  // ---- dense order
  std::vector<int> ins2dense, dense2ins;
  bool identity_order = true;

  bool topo_dirty = true;         // factors/vertices added since the last structure build
  bool host_values_fresh = true;  // host xyt newer than (or equal to) device poses
  bool theta_wrapped = false;     // every host theta is already in (-pi, pi] (they all came back from the device)
  double4* h_stage = nullptr;     // pinned staging of the pose transfers
  size_t h_stage_n = 0;
  char* h_struct[2] = {nullptr, nullptr};   // pinned staging of the structure upload, two chunks in flight
  cudaEvent_t ev_struct[2] = {nullptr, nullptr};
  bool device_values_valid = false;
  bool system_valid = false;      // H, b assembled at the current device poses

  LocalStructure S;
  std::vector<int> di, dj, dp, perm;  // dense endpoints (caller's order), dense prior rows, canonical order
  int N = 0, E = 0, P = 0;
  int n_own = 0, n_loc = 0;
  int e_count = 0;  // factors this rank is the counting side of

  // ---- device
  DevBuf<double4> d_pose, d_pose_try, d_pr_pose, d_pose_all;
  DevBuf<double4> d_pose_lin;   // two-level preconditioner: the linearisation point its basis W is built at (fixed address:
                                // d_pose / d_pose_try swap on every accepted step, the captured PCG graph must not see that)
  DevBuf<int2> d_eij, d_refs;
  DevBuf<int4> d_cw;
  DevBuf<double2> d_fdata;   // factor cells, five double2 components each (layout: rows.cuh, GS_F_LAYOUT)
  size_t fstride = gs::kSlice;
  DevBuf<int> d_pr_row, d_cbase, d_cwidth, d_rbase, d_rwidth, d_segtile, d_send_rows, d_cfirst;
  DevBuf<double> d_pr_om, d_Ho, d_Hd, d_b, d_x, d_r, d_z, d_p, d_Ap, d_Minv, d_dscale, d_part, d_segs, d_sendbuf;
  DevBuf<Scalars> d_sc;
  DevBuf<unsigned int> d_counter;
  DevBuf<gs::ClosestOut> d_closest;       // closest_keyframe: [0] result, [2 .. 2+ranks) gathered results, then per-CTA partials
  DevBuf<int> d_closest_order;            // insertion index -> dense row, when keys were not inserted in increasing order
  int closest_order_n = -1;
  Scalars* h_sc = nullptr;  // pinned mirror
  size_t hstride = 1;
  int grid_rows = 1, grid_flat3 = 1, part_stride = 1;

  cudaGraph_t pcg_graph = nullptr;
  cudaGraphExec_t pcg_exec = nullptr;

  // ---- two-level preconditioner (kernels_coarse.cuh): block-Jacobi + rigid-body coarse space over aggregates
  bool coarse_wanted = true;              // !GS_SOLVER_BLOCK_JACOBI_ONLY
  bool coarse = false;                    // in use for this graph (single rank, lambda*I damping, large enough)
  gs::Aggregation agg;
  DevBuf<int> d_agg_of, d_agg_ptr, d_agg_rows, d_agg_color, d_nbr_of_color;
  DevBuf<double> d_cvec, d_mu, d_cpart, d_R, d_G, d_E, d_chol_work, d_basis;
  DevBuf<int> d_chol_info;
  DevBuf<double2> d_agg_ref;
  DevBuf<double> d_Ework;                 // m x m scratch of coarse_inverse (L^-1)
  DevBuf<double> d_Y;                     // 9 doubles per owned row: H W_c of one colour (coarse set-up)
  DevBuf<int> d_coarse_seg_row;           // [kSegments + 1] rows of the coarse vector per fixed reduction segment (3 x seg_agg)
  int chol_lwork = 0;
  int agg_a0 = 0, agg_own = 0;            // this rank's aggregates [agg_a0, agg_a0 + agg_own)
  bool coarse_e0_valid = false;           // R (= W^T H W by columns) and G computed at the current linearisation
  double coarse_lambda = -1.0;            // the lambda d_E (= E^-1) was built for; < 0: none
  double ms_coarse_setup = 0.0;

  // ---- last optimise trace
  std::vector<double> tr_err, tr_lambda;
  std::vector<int32_t> tr_tries;
  int64_t launches = 0;
  double prep_ms[4] = {0, 0, 0, 0};   // last structure build: host structure, staging + upload, communicator setup, graph capture
  double p2p_wait[4] = {0, 0, 0, 0};  // peer-memory path: clock ticks CTA 0 waited in spmv / update / dir, waits counted

  gs_status fail(gs_status s, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    err = buf;
    return s;
  }
  int find(uint64_t key) const {
    if (keys.empty()) return -1;
    if (keys_sequential) {
      if (key < keys[0]) return -1;
      const uint64_t d = key - keys[0];
      return d < keys.size() ? (int)d : -1;
    }
    auto it = key2ins.find(key);
    return it == key2ins.end() ? -1 : it->second;
  }
};

#define GS_CUDA(g, call)                                                                         \
  do {                                                                                           \
    cudaError_t e__ = (call);                                                                    \
    if (e__ != cudaSuccess)                                                                      \
      return (g)->fail(GS_ECUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
  } while (0)

#define GS_NCCL(g, call)                                                                         \
  do {                                                                                           \
    ncclResult_t r__ = (call);                                                                   \
    if (r__ != ncclSuccess)                                                                      \
      return (g)->fail(GS_ENCCL, "%s:%d %s: %s", __FILE__, __LINE__, #call, gs::nccl().GetErrorString(r__)); \
  } while (0)

namespace {

void destroy_graph_exec(gs_graph* g) {
  if (g->pcg_exec) cudaGraphExecDestroy(g->pcg_exec);
  if (g->pcg_graph) cudaGraphDestroy(g->pcg_graph);
  g->pcg_exec = nullptr;
  g->pcg_graph = nullptr;
}

gs::ReduceCtx reduce_ctx(gs_graph* g, int counter_slot) {
  gs::ReduceCtx c;
  c.part = g->d_part.p;
  c.part_stride = g->part_stride;
  c.segs = g->d_segs.p;
  c.seg_tile = g->d_segtile.p;
  c.spr = gs::kSegments / g->world;
  c.rank = g->rank;
  c.world = g->world;
  c.counter = g->d_counter.p + counter_slot;
  c.sc = g->d_sc.p;
  return c;
}

gs::AsmView asm_view(gs_graph* g, const double4* pose) {
  gs::AsmView a;
  a.cs.base = g->d_cbase.p;
  a.cs.width = g->d_cwidth.p;
  a.c = g->d_cw.p;
  a.c_first = g->d_cfirst.p;
  a.f.ij = g->d_eij.p;
  a.f.data = g->d_fdata.p;
  a.f.fstride = g->fstride;
  a.pr.n = g->S.P_loc;
  a.pr.row = g->d_pr_row.p;
  a.pr.pose = g->d_pr_pose.p;
  a.pr.om = g->d_pr_om.p;
  a.pose = pose;
  a.Ho = g->d_Ho.p;
  a.hstride = g->hstride;
  a.Hd = g->d_Hd.p;
  a.b = g->d_b.p;
  a.n_own = g->n_own;
  return a;
}

gs::SpmvView spmv_view(gs_graph* g) {
  gs::SpmvView a;
  a.rs.base = g->d_rbase.p;
  a.rs.width = g->d_rwidth.p;
  a.refs = g->d_refs.p;
  a.Ho = g->d_Ho.p;
  a.hstride = g->hstride;
  a.Hd = g->d_Hd.p;
  a.dscale = g->prm.diagonal_damping ? g->d_dscale.p : nullptr;
  a.n_own = g->n_own;
  return a;
}

gs::PcgView pcg_view(gs_graph* g) {
  gs::PcgView a;
  a.b = g->d_b.p;
  a.Minv = g->d_Minv.p;
  a.x = g->d_x.p;
  a.r = g->d_r.p;
  a.z = g->d_z.p;
  a.p = g->pvec;
  a.Ap = g->d_Ap.p;
  a.n_own = g->n_own;
  return a;
}

bool coarse_active(const gs_graph* g) { return g->coarse && !g->prm.diagonal_damping; }

// cuSOLVER handles are expensive to create (~0.5 s with the libraries' own initialisation): one per device for the
// process, used under a lock with the calling graph's stream
struct SolverPool {
  std::mutex mu;
  cusolverDnHandle_t h[64] = {nullptr};
  cublasHandle_t hb[64] = {nullptr};
};
SolverPool& solver_pool() {
  static SolverPool p;
  return p;
}
cusolverDnHandle_t device_cusolver() {   // call with solver_pool().mu held
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  SolverPool& p = solver_pool();
  if (!p.h[dev] && cusolverDnCreate(&p.h[dev]) != CUSOLVER_STATUS_SUCCESS) p.h[dev] = nullptr;
  return p.h[dev];
}

cublasHandle_t device_cublas() {   // call with solver_pool().mu held
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  SolverPool& p = solver_pool();
  if (!p.hb[dev] && cublasCreate(&p.hb[dev]) != CUBLAS_STATUS_SUCCESS) p.hb[dev] = nullptr;
  return p.hb[dev];
}

// E (lower triangle of an SPD matrix, column-major, m x m, in Em) -> E^-1 (both triangles) in Em; X: m x m scratch.
// Cholesky by cuSOLVER, then E^-1 = L^-T L^-1 with L^-1 = trsm(L, I) and one syrk — 1.2 + 2.6 + 0.9 ms at m = 3072 on a
// B200, where cusolverDnDpotri alone takes 7.9 ms (microbenchmark recorded in profiles/r2_summary.md).  Call with
// solver_pool().mu held.  `info` receives potrf's status word (0 = positive definite) after a stream synchronisation.
gs_status coarse_inverse(gs_graph* g, int m, double* Em, double* X, int* info) {
  cusolverDnHandle_t hs = device_cusolver();
  cublasHandle_t hb = device_cublas();
  if (!hs || !hb || cusolverDnSetStream(hs, g->stream) != CUSOLVER_STATUS_SUCCESS || cublasSetStream(hb, g->stream) != CUBLAS_STATUS_SUCCESS)
    return g->fail(GS_ECUDA, "cuSOLVER / cuBLAS handle unavailable");
  if (cusolverDnDpotrf(hs, CUBLAS_FILL_MODE_LOWER, m, Em, m, g->d_chol_work.p, g->chol_lwork, g->d_chol_info.p) != CUSOLVER_STATUS_SUCCESS)
    return g->fail(GS_ECUDA, "cusolverDnDpotrf failed");
  *info = 0;
  GS_CUDA(g, cudaMemcpyAsync(info, g->d_chol_info.p, sizeof(int), cudaMemcpyDeviceToHost, g->stream));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  if (*info != 0) return GS_OK;
  const size_t mm = (size_t)m * m;
  const double one = 1.0, zero = 0.0;
  GS_CUDA(g, cudaMemsetAsync(X, 0, sizeof(double) * mm, g->stream));
  gs::k_set_diagonal<<<(unsigned)((m + gs::kThreads - 1) / gs::kThreads), gs::kThreads, 0, g->stream>>>(X, m, 1.0);
  if (cublasDtrsm(hb, CUBLAS_SIDE_LEFT, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_N, CUBLAS_DIAG_NON_UNIT, m, m, &one, Em, m, X, m) !=
          CUBLAS_STATUS_SUCCESS ||
      cublasDsyrk(hb, CUBLAS_FILL_MODE_LOWER, CUBLAS_OP_T, m, m, &one, X, m, &zero, Em, m) != CUBLAS_STATUS_SUCCESS)
    return g->fail(GS_ECUDA, "cuBLAS trsm / syrk failed");
  gs::k_symmetrize<<<(unsigned)((mm + gs::kThreads - 1) / gs::kThreads), gs::kThreads, 0, g->stream>>>(Em, m);
  g->launches += 3;
  return GS_OK;
}

gs::CoarseView coarse_view(gs_graph* g) {
  gs::CoarseView c;
  c.agg_of = g->d_agg_of.p;
  c.agg_ptr = g->d_agg_ptr.p;
  c.agg_rows = g->d_agg_rows.p;
  c.pose = g->d_pose_lin.p;
  c.ref = g->d_agg_ref.p;
  c.a0 = g->agg_a0;
  c.n_agg_own = g->agg_own;
  c.n_agg = g->agg.n_agg;
  c.n_own = g->n_own;
  return c;
}

// ---- multi-rank plumbing (no-ops on a single rank) ------------------------------------------------
// ghost slots of `vec` (width doubles per vertex) <- owner ranks' current values
// `ghost` = where the ghost rows land (nullptr: right behind the owned rows of `vec`)
gs_status halo_exchange(gs_graph* g, double* vec, int width, double* ghost = nullptr) {
  if (!ghost) ghost = vec + (size_t)g->n_own * width;
  if (g->world == 1) return GS_OK;
  const LocalStructure& S = g->S;
  const int nsend = (int)S.send_rows.size();
  if (nsend) {
    const int n = nsend * width;
    gs::k_pack<<<(n + gs::kThreads - 1) / gs::kThreads, gs::kThreads, 0, g->stream>>>(vec, g->d_send_rows.p, nsend, width,
                                                                                      g->d_sendbuf.p);
    g->launches += 1;
  }
  if (S.peers.empty()) return GS_OK;
  gs::NcclApi& nc = gs::nccl();
  GS_NCCL(g, nc.GroupStart());
  for (const gs::HaloPeer& hp : S.peers) {
    if (hp.send_cnt)
      GS_NCCL(g, nc.Send(g->d_sendbuf.p + (size_t)hp.send_off * width, (size_t)hp.send_cnt * width, ncclDouble, hp.rank,
                         g->comm, g->stream));
    if (hp.ghost_cnt)
      GS_NCCL(g, nc.Recv(ghost + (size_t)hp.ghost_off * width, (size_t)hp.ghost_cnt * width, ncclDouble, hp.rank, g->comm,
                         g->stream));
  }
  GS_NCCL(g, nc.GroupEnd());
  return GS_OK;
}

// after a reducing kernel: all-gather the segment sums, then the scalar step on every rank
gs_status finish_reduce(gs_graph* g, int nq, int op) {
  if (g->world == 1) return GS_OK;
  const int spr = gs::kSegments / g->world;
  const size_t chunk = (size_t)nq * spr;
  GS_NCCL(g, gs::nccl().AllGather(g->d_segs.p + g->rank * chunk, g->d_segs.p, chunk, ncclDouble, g->comm, g->stream));
  gs::k_scalars<<<1, 32, 0, g->stream>>>(g->d_segs.p, nq, spr, op, g->d_sc.p);
  g->launches += 1;
  return GS_OK;
}

gs_status launch_spmv(gs_graph* g, double* p, double* y, int mode) {
  if (gs_status s = halo_exchange(g, p, 3)) return s;
  gs::k_spmv<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(spmv_view(g), p, p + 3 * (size_t)g->n_own, y, reduce_ctx(g, 0),
                                                           mode);
  return finish_reduce(g, 1, mode == 0 ? gs::OP_PAP : gs::OP_DOT);
}

gs_status coarse_setup(gs_graph* g, double lambda);
gs_status coarse_gather(gs_graph* g, double* vec, int width);

template <typename T>
gs_status upload(gs_graph* g, DevBuf<T>& d, const std::vector<T>& h) {
  GS_CUDA(g, d.alloc(std::max<size_t>(h.size(), 1)));
  if (!h.empty())
    GS_CUDA(g, cudaMemcpyAsync(d.p, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, g->stream));
  g->h2d_bytes += sizeof(T) * (double)h.size();
  return GS_OK;
}

gs::P2PCtx p2p_ctx(gs_graph* g, unsigned int k_local) {
  gs::P2PCtx x;
  x.me = reinterpret_cast<gs::P2PWindow*>(g->win);
  x.win = g->d_winptr.p;
  x.world = g->world;
  x.rank = g->rank;
  x.spr = gs::kSegments / g->world;
  x.nbr_mask = g->nbr_mask;
  x.seq_base = g->d_seq.p;
  x.k_local = k_local;
  x.ticket = g->d_seq.p + 32;   // its own 128-byte line
  x.spin_limit = g->spin_limit_ticks;
  x.tile_halo = g->d_tile_halo.p;
  return x;
}

// one stage of a PCG iteration (0: SpMV + p.Ap, 1: x/r/z update + r.z, r.r, 2: direction);
// k_local = the iteration's position in the batch (peer-memory path only)
gs_status launch_pcg_stage(gs_graph* g, int stage, unsigned int k_local) {
  if (g->p2p && g->dir_otf) {
    const gs::P2PCtx x = p2p_ctx(g, k_local);
    // iteration k reads the direction of iteration k-1 from buffer k&1 and writes its own into the other one
    double* P[2] = {g->pvec, g->d_p_alt.p};
    double* prev = P[k_local & 1u];
    double* next = P[(k_local + 1u) & 1u];
    const size_t own3 = 3 * (size_t)g->n_own;
    gs::HaloPush h;
    h.row = g->d_hp_row.p;
    h.dst = g->d_hp_dst.p;
    h.n_push_tiles = g->n_push_tiles;
    h.order = g->d_hp_order.p;
    if (stage == 0) {
      gs::DirSpmvArgs d;
      d.z = g->d_z.p;
      d.zg = g->pghost;
      d.p_prev = prev;
      d.pg_prev = prev + own3;
      d.p_next = next;
      d.pg_next = next + own3;
      d.y = g->d_Ap.p;
      d.n_ghost = g->S.n_ghost;
      d.ghost_chunk = g->ghost_chunk;
      gs::k_spmvdir_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(spmv_view(g), d, reduce_ctx(g, 0), x);
    } else if (stage == 1) {
      gs::PcgView pv = pcg_view(g);
      pv.p = next;
      gs::k_updpush_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(pv, reduce_ctx(g, 1), x, h);
    }
    return GS_OK;
  }
  if (g->p2p && coarse_active(g)) {
    // two-level preconditioner over peer memory: six kernels, no NCCL call (kernels_p2p.cuh)
    const gs::P2PCtx x = p2p_ctx(g, k_local);
    const int m = 3 * g->agg.n_agg;
    if (stage == 0) {
      gs::k_spmv_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(spmv_view(g), g->pvec, g->pghost, g->d_Ap.p,
                                                                   reduce_ctx(g, 0), x);
    } else if (stage == 1) {
      gs::k_update_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(pcg_view(g), reduce_ctx(g, 1), x);
      if (g->agg_own) gs::k_restrict_p2p<<<g->agg_own, gs::kThreads, 0, g->stream>>>(coarse_view(g), g->d_r.p, g->d_sc.p, x);
      gs::k_c_unpack_p2p<<<(m + gs::kThreads - 1) / gs::kThreads, gs::kThreads, 0, g->stream>>>(x, g->agg.n_agg, g->d_cvec.p, g->d_sc.p);
      const int nrows = 3 * g->agg_own;
      if (nrows)
        gs::k_coarse_mv_p2p<<<(nrows + gs::kCoarseRows - 1) / gs::kCoarseRows, gs::kThreads, 0, g->stream>>>(
            g->d_E.p, g->d_cvec.p, m, 3 * g->agg_a0, nrows, g->d_mu.p, g->d_cpart.p, g->d_coarse_seg_row.p, g->d_counter.p + 6,
            g->d_sc.p, x);
    } else {
      gs::HaloPush h;
      h.row = g->d_hp_row.p;
      h.dst = g->d_hp_dst.p;
      h.n_push_tiles = g->n_push_tiles;
      h.order = g->d_hp_order.p;
      gs::k_dir_coarse_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(coarse_view(g), g->d_z.p, g->d_mu.p, g->pvec, g->d_sc.p, x, h);
    }
    return GS_OK;
  }
  if (g->p2p) {
    const gs::P2PCtx x = p2p_ctx(g, k_local);
    if (stage == 0) {
      gs::k_spmv_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(spmv_view(g), g->pvec, g->pghost, g->d_Ap.p,
                                                                   reduce_ctx(g, 0), x);
    } else if (stage == 1) {
      gs::k_update_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(pcg_view(g), reduce_ctx(g, 1), x);
    } else {
      gs::HaloPush h;
      h.row = g->d_hp_row.p;
      h.dst = g->d_hp_dst.p;
      h.n_push_tiles = g->n_push_tiles;
      h.order = g->d_hp_order.p;
      gs::k_dir_p2p<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(pcg_view(g), reduce_ctx(g, 1), x, h);
    }
    return GS_OK;
  }
  if (stage == 0) return launch_spmv(g, g->pvec, g->d_Ap.p, 0);
  if (coarse_active(g)) {
    // two-level preconditioner: z = M^-1 r + W E^-1 W^T r.  r.z = r.M^-1 r + c.mu is completed by k_coarse_mv, which also
    // forms beta; the corrected z is only ever needed inside the direction update
    const int m = 3 * g->agg.n_agg;
    if (stage == 1) {
      gs::k_pcg_update_staged<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(pcg_view(g), reduce_ctx(g, 1), gs::OP_UPDATE_2L);
      if (gs_status s = finish_reduce(g, 2, gs::OP_UPDATE_2L)) return s;
      if (g->agg_own)
        gs::k_restrict<<<g->agg_own, gs::kThreads, 0, g->stream>>>(coarse_view(g), g->d_r.p, g->d_sc.p, 1, g->d_cvec.p, 3, 0);
      if (gs_status s = coarse_gather(g, g->d_cvec.p, 3)) return s;
      gs::k_coarse_mv<<<(m + gs::kCoarseRows - 1) / gs::kCoarseRows, gs::kThreads, 0, g->stream>>>(
          g->d_E.p, g->d_cvec.p, m, g->d_mu.p, g->d_cpart.p, g->d_coarse_seg_row.p, g->d_counter.p + 6, g->d_sc.p, 0);
    } else {
      gs::k_dir_coarse<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(coarse_view(g), g->d_z.p, g->d_mu.p, g->pvec, g->d_sc.p);
    }
    return GS_OK;
  }
  if (stage == 1) {
    gs::k_pcg_update_staged<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(pcg_view(g), reduce_ctx(g, 1), gs::OP_UPDATE);
    return finish_reduce(g, 2, gs::OP_UPDATE);
  }
  gs::k_pcg_dir<<<g->grid_flat3, gs::kThreads, 0, g->stream>>>(g->d_z.p, g->pvec, g->d_sc.p, 3 * g->n_own);
  return GS_OK;
}

gs_status launch_pcg_iteration(gs_graph* g, unsigned int k_local) {
  for (int stage = 0; stage < 3; ++stage)
    if (gs_status s = launch_pcg_stage(g, stage, k_local)) return s;
  return GS_OK;
}

// the persistent small-graph kernel: `n` iterations in one cooperative launch
gs_status launch_pcg_fused(gs_graph* g, int n) {
  gs::SpmvView sv = spmv_view(g);
  gs::PcgView pv = pcg_view(g);
  double* Ap = g->d_Ap.p;
  double* p_alt = g->d_p_alt.p;
  double* part = g->d_part.p;
  int stride = g->part_stride;
  const int* seg_tile = g->d_segtile.p;
  Scalars* sc = g->d_sc.p;
  gs::GridBarrier* bar = g->d_bar.p;
  void* args[] = {&sv, &pv, &Ap, &p_alt, &part, &stride, &seg_tile, &sc, &bar, &n};
  GS_CUDA(g, cudaLaunchCooperativeKernel((const void*)gs::k_pcg_fused, dim3((unsigned)g->grid_rows), dim3(gs::kThreads), args, 0,
                                         g->stream));
  return GS_OK;
}

// a batch of iterations; the peer-memory path then moves the sequence base
gs_status launch_pcg_batch(gs_graph* g, int n) {
  for (int it = 0; it < n; ++it)
    if (gs_status s = launch_pcg_iteration(g, (unsigned int)it)) return s;
  if (g->p2p) gs::k_p2p_advance<<<1, 32, 0, g->stream>>>(g->d_seq.p, (unsigned int)n);
  return GS_OK;
}

gs_status capture_pcg_graph(gs_graph* g) {
  destroy_graph_exec(g);
  if (!g->use_graph) return GS_OK;
  const int64_t keep = g->launches;
  GS_CUDA(g, cudaStreamBeginCapture(g->stream, cudaStreamCaptureModeThreadLocal));
  const gs_status rc = launch_pcg_batch(g, kPcgBatch);
  cudaError_t ce = cudaStreamEndCapture(g->stream, &g->pcg_graph);
  g->launches = keep;
  if (rc != GS_OK) return rc;
  GS_CUDA(g, ce);
  GS_CUDA(g, cudaGraphInstantiate(&g->pcg_exec, g->pcg_graph, 0));
  return GS_OK;
}

// ---- peer-memory window management ---------------------------------------------------------------
struct XchgRecord {            // what every rank tells every other rank at prepare time (128 bytes)
  cudaIpcMemHandle_t handle;   // 64
  long long offset;            // window pointer - allocation base
  int generation;              // bumps when the window is reallocated
  int will_realloc;
  int n_own;
  int ghost_off[gs::kMaxRanks];  // [source rank] first ghost slot filled by that rank
  int pad[3];
};
static_assert(sizeof(XchgRecord) == 128, "XchgRecord must be 128 bytes");

long long allocation_offset(void* ptr) {
  // cudaIpcOpenMemHandle maps the whole underlying allocation: find our offset inside it
  typedef int (*Fn)(unsigned long long*, size_t*, unsigned long long);
  static Fn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    if (void* h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL)) fn = (Fn)dlsym(h, "cuMemGetAddressRange_v2");
  }
  if (!fn) return -1;   // cannot be resolved: the caller fails instead of letting peers write at a wrong address
  unsigned long long base = 0;
  size_t size = 0;
  if (fn(&base, &size, (unsigned long long)(uintptr_t)ptr) != 0) return -1;
  return (long long)((unsigned long long)(uintptr_t)ptr - base);
}

// all-gather of one 128-byte record per rank (NCCL, through a small device buffer)
gs_status exchange_records(gs_graph* g, const XchgRecord& mine, std::vector<XchgRecord>& all) {
  const size_t rec = sizeof(XchgRecord);
  GS_CUDA(g, g->d_xchg.alloc(rec * gs::kMaxRanks));
  GS_CUDA(g, cudaMemcpyAsync(g->d_xchg.p + rec * g->rank, &mine, rec, cudaMemcpyHostToDevice, g->stream));
  GS_NCCL(g, gs::nccl().AllGather(g->d_xchg.p + rec * g->rank, g->d_xchg.p, rec, ncclChar, g->comm, g->stream));
  all.resize(g->world);
  GS_CUDA(g, cudaMemcpyAsync(all.data(), g->d_xchg.p, rec * g->world, cudaMemcpyDeviceToHost, g->stream));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  return GS_OK;
}

void close_peer(gs_graph* g, int q) {
  if (g->peer_base[q]) cudaIpcCloseMemHandle(g->peer_base[q]);
  g->peer_base[q] = nullptr;
  g->peer_win[q] = nullptr;
}

// (re)build the window, exchange handles, map the peers, and lay out the halo pushes.  Collective.
gs_status prepare_p2p(gs_graph* g) {
  const LocalStructure& S = g->S;
  // window = header | landing zone of the coarse vector c (two-level preconditioner; 2 parities x 6 n_agg flagged words,
  // the same size on every rank) | ghost rows of p
  g->win_ghost_off = gs::kWindowHeaderBytes + (g->coarse ? ((2 * 6 * (size_t)g->agg.n_agg * 8 + 255) / 256) * 256 : 0);
  const size_t need = g->win_ghost_off + sizeof(double) * 3 * (size_t)std::max(S.n_ghost, 1);
  XchgRecord mine;
  std::memset(&mine, 0, sizeof mine);
  mine.will_realloc = (need > g->win_bytes) ? 1 : 0;
  std::vector<XchgRecord> all;
  // round 1: who reallocates?  close those mappings before anybody frees
  if (gs_status s = exchange_records(g, mine, all)) return s;
  for (int q = 0; q < g->world; ++q)
    if (q != g->rank && all[q].will_realloc) close_peer(g, q);
  if (gs_status s = exchange_records(g, mine, all)) return s;  // barrier: everybody has closed
  if (mine.will_realloc) {
    if (g->win) GS_CUDA(g, cudaFree(g->win));
    g->win = nullptr;
    const size_t two_mb = (size_t)2 << 20;
    g->win_bytes = ((need + need / 4 + two_mb - 1) / two_mb) * two_mb;
    GS_CUDA(g, cudaMalloc(&g->win, g->win_bytes));
    GS_CUDA(g, cudaMemsetAsync(g->win, 0, g->win_bytes, g->stream));
    g->win_generation += 1;
    // the sequence numbering restarts with the fresh (zeroed) flags on every rank at once?  No: only this
    // rank's flags are new, so keep counting — stamps only ever need to exceed what a waiter asks for.
  }
  g->pghost = reinterpret_cast<double*>(g->win + g->win_ghost_off);
  if (g->win_ghost_off > gs::kWindowHeaderBytes)   // stamps of an earlier layout must not survive in the landing zone
    GS_CUDA(g, cudaMemsetAsync(g->win + gs::kWindowHeaderBytes, 0, g->win_ghost_off - gs::kWindowHeaderBytes, g->stream));
  GS_CUDA(g, cudaMemsetAsync(g->pghost, 0, sizeof(double) * 3 * (size_t)std::max(S.n_ghost, 1), g->stream));
  GS_CUDA(g, cudaIpcGetMemHandle(&mine.handle, g->win));
  mine.offset = allocation_offset(g->win);
  if (mine.offset < 0)
    return g->fail(GS_ECUDA, "cuMemGetAddressRange unavailable: the peer-memory window cannot be located inside its allocation "
                             "(create the graph with GS_SOLVER_NCCL_PER_OP)");
  mine.generation = g->win_generation;
  mine.n_own = g->n_own;
  for (const gs::HaloPeer& hp : S.peers) mine.ghost_off[hp.rank] = hp.ghost_off;
  if (gs_status s = exchange_records(g, mine, all)) return s;
  for (int q = 0; q < g->world; ++q) {
    if (q == g->rank) continue;
    if (g->peer_base[q] && g->peer_generation[q] == all[q].generation) continue;
    close_peer(g, q);
    void* base = nullptr;
    GS_CUDA(g, cudaIpcOpenMemHandle(&base, all[q].handle, cudaIpcMemLazyEnablePeerAccess));
    g->peer_base[q] = base;
    g->peer_win[q] = static_cast<char*>(base) + all[q].offset;
    g->peer_generation[q] = all[q].generation;
  }
  // halo pushes: my k-th send row for rank q lands in q's ghost slot ghost_off_q[me] + k
  struct Entry {
    int row;
    double* dst;
  };
  std::vector<Entry> ent;
  g->nbr_mask = 0;
  for (const gs::HaloPeer& hp : S.peers) {
    g->nbr_mask |= 1u << hp.rank;
    double* peer_p = reinterpret_cast<double*>(g->peer_win[hp.rank] + g->win_ghost_off);
    const size_t slot0 = (size_t)all[hp.rank].ghost_off[g->rank];
    for (int k = 0; k < hp.send_cnt; ++k) ent.push_back({S.send_rows[hp.send_off + k], peer_p + 3 * (slot0 + k)});
  }
  std::stable_sort(ent.begin(), ent.end(), [](const Entry& a, const Entry& b) { return a.row < b.row; });
  const int nt = g->grid_rows;
  std::vector<int> tile_ptr(nt + 1, 0), rows(ent.size());
  std::vector<double*> dsts(ent.size());
  for (size_t e = 0; e < ent.size(); ++e) {
    rows[e] = ent[e].row;
    dsts[e] = ent[e].dst;
    tile_ptr[ent[e].row / gs::kTile + 1]++;
  }
  for (int t = 0; t < nt; ++t) tile_ptr[t + 1] += tile_ptr[t];
  std::vector<int4> order;
  order.reserve(nt);
  for (int t = 0; t < nt; ++t)
    if (tile_ptr[t + 1] > tile_ptr[t]) order.push_back(make_int4(t, tile_ptr[t], tile_ptr[t + 1], 0));
  g->n_push_tiles = (int)order.size();
  for (int t = 0; t < nt; ++t)
    if (tile_ptr[t + 1] == tile_ptr[t]) order.push_back(make_int4(t, 0, 0, 0));
  if (gs_status s = upload(g, g->d_hp_order, order)) return s;
  if (gs_status s = upload(g, g->d_hp_row, rows)) return s;
  if (gs_status s = upload(g, g->d_hp_dst, dsts)) return s;
  // Two kernels per iteration pay a second gather per column in the product and save the direction kernel:
  // measured on C3, 1.39 -> 1.31 s per step at 2 GPUs and 0.92 -> 0.87 s at 4, but 0.54 -> 0.6 s at 8 (125k rows
  // per rank), so the choice goes by rows per rank (the same number on every rank).
  GS_CUDA(g, g->d_p_alt.alloc(3 * (size_t)std::max(S.n_own + S.n_ghost, 1) + 3));
  std::vector<int> tile_halo(nt, 0);
  for (int r = 0; r < g->n_own; ++r) {
    const int w = S.rsell.width[r / gs::kSlice];
    for (int k = 0; k < w; ++k)
      if (S.r_col[S.rsell.pos(r, k)] >= g->n_own) {
        tile_halo[r / gs::kTile] = 1;
        break;
      }
  }
  int n_halo_tiles = 0;
  for (int t = 0; t < nt; ++t)
    if (tile_halo[t]) tile_halo[t] = ++n_halo_tiles;   // ordinal: which share of the ghost rows the tile advances
  g->ghost_chunk = n_halo_tiles ? (3 * S.n_ghost + n_halo_tiles - 1) / n_halo_tiles : 0;
  if (gs_status s = upload(g, g->d_tile_halo, tile_halo)) return s;
  std::vector<gs::P2PWindow*> wins(gs::kMaxRanks, nullptr);
  for (int q = 0; q < g->world; ++q) wins[q] = reinterpret_cast<gs::P2PWindow*>(q == g->rank ? g->win : g->peer_win[q]);
  if (gs_status s = upload(g, g->d_winptr, wins)) return s;
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  return GS_OK;
}

// sort keys -> dense order; identity when keys were inserted in increasing order (the reference's ids)
void build_dense_order(gs_graph* g) {
  const int n = (int)g->keys.size();
  g->identity_order = true;
  for (int k = 1; k < n; ++k)
    if (g->keys[k] <= g->keys[k - 1]) {
      g->identity_order = false;
      break;
    }
  g->dense2ins.resize(n);
  g->ins2dense.resize(n);
  for (int k = 0; k < n; ++k) g->dense2ins[k] = k;
  if (!g->identity_order)
    std::sort(g->dense2ins.begin(), g->dense2ins.end(), [&](int a, int b) { return g->keys[a] < g->keys[b]; });
  for (int k = 0; k < n; ++k) g->ins2dense[g->dense2ins[k]] = k;
}

gs_status ensure_stage(gs_graph* g, size_t n) {
  if (n <= g->h_stage_n) return GS_OK;
  if (g->h_stage) cudaFreeHost(g->h_stage);
  g->h_stage = nullptr;
  g->h_stage_n = 0;
  const size_t want = n + n / 8 + 1024;
  GS_CUDA(g, cudaMallocHost(&g->h_stage, sizeof(double4) * want));
  g->h_stage_n = want;
  return GS_OK;
}

// Pinned staging of the structure upload.  The factor cells and the contribution / reference words (90 % of the
// structure's bytes) are written by several host threads into one of two fixed pinned chunks while the other one is
// on its way to the device: no pageable vectors (a page fault per 4 KB and a synchronous staged copy on every
// rebuild of an incremental graph), and no pinned allocation that grows with the graph (pinning 800 MB for the
// one-shot upload of a million poses costs more than the copy).
constexpr size_t kStructChunkCells = 131072;                  // cells per chunk, a multiple of kSlice
constexpr size_t kStructChunkBytes = kStructChunkCells * 96;  // 8 (pair) + 80 (factor) per factor cell is the largest item

gs_status ensure_struct_stage(gs_graph* g) {
  for (int b = 0; b < 2; ++b) {
    if (!g->h_struct[b]) GS_CUDA(g, cudaMallocHost(&g->h_struct[b], kStructChunkBytes));
    if (!g->ev_struct[b]) GS_CUDA(g, cudaEventCreateWithFlags(&g->ev_struct[b], cudaEventDisableTiming));
  }
  return GS_OK;
}

// device values -> host xyt.  Multi-rank: collective (every rank broadcasts its owned range).
gs_status download_values(gs_graph* g) {
  if (!g->device_values_valid || g->host_values_fresh) return GS_OK;
  const int n = g->N;  // vertices present at the last upload
  if (gs_status s = ensure_stage(g, (size_t)n)) return s;
  double4* tmp = g->h_stage;
  const double4* src = g->d_pose.p;
  if (g->world > 1) {
    gs::NcclApi& nc = gs::nccl();
    GS_NCCL(g, nc.GroupStart());
    for (int q = 0; q < g->world; ++q) {
      const int b = g->S.part.row_begin(q), e = g->S.part.row_end(q);
      if (e > b)
        GS_NCCL(g, nc.Broadcast(g->d_pose.p, g->d_pose_all.p + b, (size_t)(e - b) * 4, ncclDouble, q, g->comm, g->stream));
    }
    GS_NCCL(g, nc.GroupEnd());
    src = g->d_pose_all.p;
  }
  GS_CUDA(g, cudaMemcpyAsync(tmp, src, sizeof(double4) * n, cudaMemcpyDeviceToHost, g->stream));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  g->d2h_bytes += sizeof(double4) * (double)n;
  gs::parallel_for((size_t)n, [g, tmp](size_t b, size_t e) {
    for (size_t d = b; d < e; ++d) {
      const int k = g->dense2ins[d];
      g->xyt[3 * k] = tmp[d].x;
      g->xyt[3 * k + 1] = tmp[d].y;
      g->xyt[3 * k + 2] = std::atan2(tmp[d].w, tmp[d].z);
    }
  });
  g->host_values_fresh = true;
  g->theta_wrapped = (size_t)n == g->keys.size();
  return GS_OK;
}

// host xyt -> device poses of this rank (owned rows, then ghosts)
gs_status upload_values(gs_graph* g) {
  const int n = g->n_loc;
  if (gs_status s = ensure_stage(g, (size_t)std::max(n, 1))) return s;
  double4* tmp = g->h_stage;
  gs::parallel_for((size_t)n, [g, tmp](size_t b, size_t e) {
    for (size_t l = b; l < e; ++l) {
      const int d = (int)l < g->n_own ? g->S.v0 + (int)l : g->S.ghosts[l - g->n_own];
      const int k = g->dense2ins[d];
      tmp[l] = make_double4(g->xyt[3 * k], g->xyt[3 * k + 1], std::cos(g->xyt[3 * k + 2]), std::sin(g->xyt[3 * k + 2]));
    }
  });
  if (n) GS_CUDA(g, cudaMemcpyAsync(g->d_pose.p, tmp, sizeof(double4) * n, cudaMemcpyHostToDevice, g->stream));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  g->h2d_bytes += sizeof(double4) * (double)n;
  g->device_values_valid = true;
  g->system_valid = false;
  return GS_OK;
}

gs_status prepare(gs_graph* g, double* ms_out) {
  if (ms_out) *ms_out = 0.0;
  NvtxRange nvtx_range(g->topo_dirty ? "gs:prepare (structure build)" : "gs:prepare");
  if (!g->topo_dirty) {
    if (g->host_values_fresh && !g->device_values_valid) return upload_values(g);
    return GS_OK;
  }
  const double t0 = now_ms();
  // values may live on the device (warm start, graph.cpp:130): bring them home before re-indexing
  if (gs_status s = download_values(g)) return s;
  if (g->keys.empty()) return g->fail(GS_EINVAL, "graph has no vertices");
  build_dense_order(g);
  g->closest_order_n = -1;
  const int N = (int)g->keys.size(), E = (int)g->e1.size(), P = (int)g->pr_v.size();
  g->di.resize(E);
  g->dj.resize(E);
  g->dp.resize(P);
  for (int k = 0; k < E; ++k) {
    g->di[k] = g->ins2dense[g->e1[k]];
    g->dj[k] = g->ins2dense[g->e2[k]];
  }
  for (int k = 0; k < P; ++k) g->dp[k] = g->ins2dense[g->pr_v[k]];
  gs::canonical_order(N, E, g->di.data(), g->dj.data(), g->perm);
  gs::build_local_structure(N, E, g->di.data(), g->dj.data(), g->perm, P, g->dp.data(), g->world, g->rank, g->S);
  const LocalStructure& S = g->S;
  // two-level preconditioner: aggregates of the block rows (small graphs keep block-Jacobi: persistent kernel on one GPU)
  // (from 1,024 poses: C1, 3,500 poses, converges in 115 ms with it and in 208 ms with the persistent block-Jacobi kernel)
  g->coarse = g->coarse_wanted && N >= kCoarseMinPoses;
  if (g->coarse) {
    std::vector<int32_t> rp, ci;
    gs::global_pattern(N, E, g->di.data(), g->dj.data(), rp, ci);
    // about sqrt(N) aggregates of about sqrt(N) poses: the dense coarse product (9 n_agg^2 doubles per PCG iteration)
    // stays a fraction of the fine-level traffic (~600 N bytes) at every size; at most 1024 aggregates
    int target = g->prm.coarse_aggregate_size > 0 ? g->prm.coarse_aggregate_size
                                                  : std::max(48, (int)std::ceil(std::sqrt((double)N)));
    target = std::max(target, (N + 4095) / 4096);   // the dense coarse inverse is (3 n_agg)^2 doubles: at most 1.2 GB
    gs::build_aggregation(N, rp, ci, target, g->agg);
  }
  const double t_host = now_ms();
  g->prep_ms[0] = t_host - t0;
  g->N = N;
  g->E = E;
  g->P = P;
  g->n_own = S.n_own;
  g->n_loc = S.n_own + S.n_ghost;
  // factor arrays in storage (SELL) order and the contribution / reference words: through the pinned chunks, below
  const int FS = S.csell.slots;
  g->fstride = gs::factor_stride(FS);
  const size_t n_cw = S.c_code.size(), n_refs = S.r_col.size();
  g->e_count = 0;
  for (uint32_t code : S.c_code)
    if (code != gs::kPadCode && (code & 4u) && (code & 3u) < 2) g->e_count++;
  std::vector<double4> prp(S.P_loc);
  std::vector<double> prom(6 * (size_t)S.P_loc);
  for (int k = 0; k < S.P_loc; ++k) {
    const int s = S.pr_src[k];
    prp[k] = make_double4(g->pr_xyt[3 * s], g->pr_xyt[3 * s + 1], std::cos(g->pr_xyt[3 * s + 2]), std::sin(g->pr_xyt[3 * s + 2]));
    for (int m = 0; m < 6; ++m) prom[6 * (size_t)k + m] = g->pr_om[6 * (size_t)s + m];
  }
  std::vector<int> segtile(S.seg_tile, S.seg_tile + gs::kSegments + 1);
  GS_CUDA(g, g->d_eij.alloc(g->fstride));
  GS_CUDA(g, g->d_fdata.alloc(5 * g->fstride));
  GS_CUDA(g, g->d_cw.alloc(std::max<size_t>(n_cw, 1)));
  GS_CUDA(g, g->d_refs.alloc(std::max<size_t>(n_refs, 1)));
  {
    if (gs_status s = ensure_struct_stage(g)) return s;
    const size_t fstride = g->fstride;
    const double* ez = g->ez.data();
    const double* eom = g->eom.data();
    int turn = 0;
    // one chunk: wait until the buffer's previous copy has left, fill it on the host threads, send it
    auto chunk = [&](auto&& fill_and_send) -> gs_status {
      const int b = turn++ & 1;
      GS_CUDA(g, cudaEventSynchronize(g->ev_struct[b]));
      if (gs_status s = fill_and_send(g->h_struct[b])) return s;
      GS_CUDA(g, cudaEventRecord(g->ev_struct[b], g->stream));
      return GS_OK;
    };
    for (size_t q0 = 0; q0 < fstride; q0 += kStructChunkCells) {          // factor cells [q0, q1): pair + five components
      const size_t q1 = std::min(fstride, q0 + kStructChunkCells), nc = q1 - q0;
      if (gs_status s = chunk([&](char* buf) -> gs_status {
            int2* c_ij = reinterpret_cast<int2*>(buf);
            double2* c_data = reinterpret_cast<double2*>(buf + sizeof(int2) * kStructChunkCells);
            // the chunk is laid out like an array of nc cells (q0 is a multiple of kSlice): slice-major chunks are one
            // contiguous piece of the device array, plane-major ones five
            gs::parallel_for(nc, [&S, c_ij, c_data, nc, q0, FS, ez, eom](size_t b, size_t e) {
              for (size_t l = b; l < e; ++l) {
                const size_t q = q0 + l;
                const int src = q < (size_t)FS ? S.f_src[q] : -1;
                if (src < 0) gs::factor_cell_default(c_ij, c_data, nc, (int)l);
                else
                  gs::factor_cell_set(c_ij, c_data, nc, (int)l, S.f_i[q] | (S.f_count[q] ? 0 : gs::kNoCountBit), S.f_j[q],
                                      ez + 3 * (size_t)src, eom + 6 * (size_t)src);
              }
            }, 8192);
            GS_CUDA(g, cudaMemcpyAsync(g->d_eij.p + q0, c_ij, sizeof(int2) * nc, cudaMemcpyHostToDevice, g->stream));
            if (gs::f_step(fstride) == (size_t)gs::kSlice) {
              GS_CUDA(g, cudaMemcpyAsync(g->d_fdata.p + gs::f_base((int)q0), c_data, sizeof(double2) * 5 * nc, cudaMemcpyHostToDevice,
                                         g->stream));
            } else {
              for (int k = 0; k < 5; ++k)
                GS_CUDA(g, cudaMemcpyAsync(g->d_fdata.p + gs::f_base((int)q0) + k * gs::f_step(fstride), c_data + k * gs::f_step(nc),
                                           sizeof(double2) * nc, cudaMemcpyHostToDevice, g->stream));
            }
            return GS_OK;
          }))
        return s;
    }
    for (size_t q0 = 0; q0 < n_cw; q0 += kStructChunkCells) {             // contribution words
      const size_t nc = std::min(n_cw, q0 + kStructChunkCells) - q0;
      if (gs_status s = chunk([&](char* buf) -> gs_status {
            int4* c = reinterpret_cast<int4*>(buf);
            gs::parallel_for(nc, [&S, c, q0](size_t b, size_t e) {
              for (size_t l = b; l < e; ++l) c[l] = make_int4((int)S.c_code[q0 + l], S.c_blk[q0 + l], S.c_col[q0 + l], 0);
            }, 16384);
            GS_CUDA(g, cudaMemcpyAsync(g->d_cw.p + q0, c, sizeof(int4) * nc, cudaMemcpyHostToDevice, g->stream));
            return GS_OK;
          }))
        return s;
    }
    for (size_t q0 = 0; q0 < n_refs; q0 += kStructChunkCells) {           // reference words
      const size_t nc = std::min(n_refs, q0 + kStructChunkCells) - q0;
      if (gs_status s = chunk([&](char* buf) -> gs_status {
            int2* c = reinterpret_cast<int2*>(buf);
            gs::parallel_for(nc, [&S, c, q0](size_t b, size_t e) {
              for (size_t l = b; l < e; ++l) c[l] = make_int2(S.r_col[q0 + l], S.r_blk[q0 + l]);
            }, 16384);
            GS_CUDA(g, cudaMemcpyAsync(g->d_refs.p + q0, c, sizeof(int2) * nc, cudaMemcpyHostToDevice, g->stream));
            return GS_OK;
          }))
        return s;
    }
    g->h2d_bytes += (double)(sizeof(int2) * fstride + sizeof(double2) * 5 * fstride + sizeof(int4) * n_cw + sizeof(int2) * n_refs);
  }
  if (gs_status s = upload(g, g->d_pr_row, S.pr_row)) return s;
  if (gs_status s = upload(g, g->d_pr_pose, prp)) return s;
  if (gs_status s = upload(g, g->d_pr_om, prom)) return s;
  if (gs_status s = upload(g, g->d_cbase, S.csell.base)) return s;
  if (gs_status s = upload(g, g->d_cwidth, S.csell.width)) return s;
  if (gs_status s = upload(g, g->d_rbase, S.rsell.base)) return s;
  if (gs_status s = upload(g, g->d_rwidth, S.rsell.width)) return s;
  if (gs_status s = upload(g, g->d_segtile, segtile)) return s;
  if (gs_status s = upload(g, g->d_cfirst, S.c_first)) return s;
  if (gs_status s = upload(g, g->d_send_rows, S.send_rows)) return s;
  const size_t n3 = 3 * (size_t)std::max(S.n_own, 1), nl3 = 3 * (size_t)std::max(g->n_loc, 1);
  g->hstride = (size_t)std::max(S.rsell.slots, gs::kSlice);
  GS_CUDA(g, g->d_pose.alloc(std::max(g->n_loc, 1)));
  GS_CUDA(g, g->d_pose_try.alloc(std::max(g->n_loc, 1)));
  if (g->world > 1) GS_CUDA(g, g->d_pose_all.alloc(N));
  GS_CUDA(g, g->d_Ho.alloc(9 * g->hstride));
  GS_CUDA(g, cudaMemsetAsync(g->d_Ho.p, 0, sizeof(double) * 9 * g->hstride, g->stream));  // padding slots stay 0
  GS_CUDA(g, g->d_Hd.alloc(2 * n3));
  GS_CUDA(g, g->d_Minv.alloc(2 * n3));
  GS_CUDA(g, g->d_b.alloc(n3));
  GS_CUDA(g, g->d_x.alloc(nl3));
  GS_CUDA(g, g->d_r.alloc(n3));
  GS_CUDA(g, g->d_z.alloc(n3));
  GS_CUDA(g, g->d_p.alloc(nl3));
  g->pvec = g->d_p.p;
  g->pghost = g->d_p.p + 3 * (size_t)S.n_own;
  GS_CUDA(g, g->d_Ap.alloc(n3));
  GS_CUDA(g, g->d_dscale.alloc(n3));
  GS_CUDA(g, g->d_sendbuf.alloc(4 * std::max<size_t>(S.send_rows.size(), 1)));
  GS_CUDA(g, cudaMemsetAsync(g->d_x.p, 0, sizeof(double) * nl3, g->stream));
  GS_CUDA(g, cudaMemsetAsync(g->pvec, 0, sizeof(double) * nl3, g->stream));
  g->grid_rows = std::max(S.ntiles, 1);
  g->grid_flat3 = (int)((3 * (size_t)S.n_own + gs::kThreads * gs::kDirPerThread - 1) / (gs::kThreads * gs::kDirPerThread));
  if (g->grid_flat3 < 1) g->grid_flat3 = 1;
  g->part_stride = g->grid_rows;
  GS_CUDA(g, g->d_part.alloc(3 * (size_t)g->part_stride));
  g->fused = g->fused_wanted && g->world == 1 && g->fused_capacity > 0 && g->grid_rows <= g->fused_capacity && !g->coarse;
  if (g->coarse) {
    const gs::Aggregation& A = g->agg;
    const size_t m = 3 * (size_t)A.n_agg;
    // this rank's aggregates: those of its segments; lists in local row indices
    const int spr = gs::kSegments / g->world;
    g->agg_a0 = A.seg_agg[g->rank * spr];
    g->agg_own = A.seg_agg[(g->rank + 1) * spr] - g->agg_a0;
    std::vector<int> l_of(A.agg_of.begin() + S.v0, A.agg_of.begin() + S.v1);   // owned rows, then ghost rows
    for (int gr : S.ghosts) l_of.push_back(A.agg_of[gr]);
    std::vector<int> l_ptr(g->agg_own + 1), l_rows;
    const int q0 = A.ptr[g->agg_a0];
    for (int a = 0; a <= g->agg_own; ++a) l_ptr[a] = A.ptr[g->agg_a0 + a] - q0;
    l_rows.assign(A.rows.begin() + q0, A.rows.begin() + A.ptr[g->agg_a0 + g->agg_own]);
    for (int& r : l_rows) r -= S.v0;
    if (gs_status s = upload(g, g->d_agg_of, l_of)) return s;
    if (gs_status s = upload(g, g->d_agg_ptr, l_ptr)) return s;
    if (gs_status s = upload(g, g->d_agg_rows, l_rows)) return s;
    if (gs_status s = upload(g, g->d_agg_color, A.color)) return s;
    if (gs_status s = upload(g, g->d_nbr_of_color, A.nbr_of_color)) return s;
    GS_CUDA(g, g->d_cvec.alloc(m));
    GS_CUDA(g, g->d_mu.alloc(m));
    GS_CUDA(g, g->d_cpart.alloc(m + 2));   // per-row products mu_i c_i
    {
      std::vector<int> seg_row(gs::kSegments + 1);
      for (int k = 0; k <= gs::kSegments; ++k) seg_row[k] = 3 * A.seg_agg[k];
      if (gs_status s = upload(g, g->d_coarse_seg_row, seg_row)) return s;
    }
    GS_CUDA(g, g->d_R.alloc((size_t)A.n_colors * 3 * m));
    GS_CUDA(g, cudaMemsetAsync(g->d_R.p, 0, sizeof(double) * (size_t)A.n_colors * 3 * m, g->stream));
    GS_CUDA(g, g->d_G.alloc(3 * m));
    GS_CUDA(g, g->d_E.alloc(m * m));
    GS_CUDA(g, g->d_basis.alloc(nl3));
    GS_CUDA(g, g->d_Y.alloc(3 * n3));
    GS_CUDA(g, g->d_pose_lin.alloc(std::max(g->n_loc, 1)));
    GS_CUDA(g, g->d_chol_info.alloc(1));
    GS_CUDA(g, g->d_agg_ref.alloc((size_t)A.n_agg));
    GS_CUDA(g, g->d_Ework.alloc(m * m));
    std::lock_guard<std::mutex> lock(solver_pool().mu);
    cusolverDnHandle_t hs = device_cusolver();
    if (!hs) return g->fail(GS_ECUDA, "cusolverDnCreate failed");
    int l1 = 0;
    if (cusolverDnDpotrf_bufferSize(hs, CUBLAS_FILL_MODE_LOWER, (int)m, g->d_E.p, (int)m, &l1) != CUSOLVER_STATUS_SUCCESS)
      return g->fail(GS_ECUDA, "cusolver workspace query failed");
    g->chol_lwork = l1;
    GS_CUDA(g, g->d_chol_work.alloc((size_t)std::max(g->chol_lwork, 1)));
    // one Cholesky + inverse of an identity of this size now: cuSOLVER and cuBLAS load their kernels on first use
    // (hundreds of milliseconds), which belongs to the structure phase, not to the first solve
    {
      int dev = 0;
      GS_CUDA(g, cudaGetDevice(&dev));
      static bool warmed[64] = {false};   // under solver_pool().mu
      if (dev >= 0 && dev < 64 && !warmed[dev]) {
        GS_CUDA(g, cudaMemsetAsync(g->d_E.p, 0, sizeof(double) * m * m, g->stream));
        gs::k_set_diagonal<<<(unsigned)((m + gs::kThreads - 1) / gs::kThreads), gs::kThreads, 0, g->stream>>>(g->d_E.p, (int)m, 1.0);
        int info = 0;
        if (gs_status s = coarse_inverse(g, (int)m, g->d_E.p, g->d_Ework.p, &info)) return s;
        GS_CUDA(g, cudaStreamSynchronize(g->stream));
        if (info != 0) return g->fail(GS_ECUDA, "cuSOLVER warm-up failed");
        warmed[dev] = true;
      }
    }
  }
  g->coarse_e0_valid = false;
  g->coarse_lambda = -1.0;
  {
    // direction on the fly (two kernels per iteration) or as a kernel of its own (three)?  Forming it inside the
    // product pays a second gather per column and saves a kernel plus, on several GPUs, a cross-GPU signal: measured
    // on C3, 1.39 -> 1.31 s per step at 2 GPUs and 0.92 -> 0.87 s at 4, but 0.54 -> 0.6 s at 8 (125k rows per rank),
    // so the choice goes by rows per rank (the same number on every rank).  The NCCL-per-exchange path keeps three.
    const int f = g->prm.solver_flags;
    // (one GPU keeps three: forming the direction inside the product was measured at 161.9 us per iteration against
    //  160.8 us on C3, and 218 us against 157 us on C4; profiles/r2_variant_probe.md)
    const bool can = g->world > 1 && g->p2p_wanted;
    if (!can || (f & GS_SOLVER_DIR_KERNEL)) g->dir_otf = false;
    else if (f & GS_SOLVER_DIR_ON_THE_FLY) g->dir_otf = true;
    else g->dir_otf = g->world > 1 ? (N / g->world >= 200000) : false;
    if (g->coarse) g->dir_otf = false;   // the two-level iteration folds the prolongation into its own direction kernel
  }
  if (g->fused) GS_CUDA(g, g->d_p_alt.alloc(3 * (size_t)std::max(g->n_own, 1) + 3));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));  // staging vectors go out of scope
  g->topo_dirty = false;
  g->device_values_valid = false;
  if (gs_status s = upload_values(g)) return s;
  const double t_up = now_ms();
  g->prep_ms[1] = t_up - t_host;
  if (g->world > 1) {
    if (g->p2p_wanted) {
      if (gs_status s = prepare_p2p(g)) return s;
      g->p2p = true;
    }
    // establish every NCCL connection outside stream capture: one halo exchange per width, one all-gather
    if (gs_status s = halo_exchange(g, g->pvec, 3)) return s;
    if (gs_status s = halo_exchange(g, reinterpret_cast<double*>(g->d_pose_try.p), 4)) return s;
    const size_t chunk = 2 * (size_t)(gs::kSegments / g->world);
    GS_NCCL(g, gs::nccl().AllGather(g->d_segs.p + g->rank * chunk, g->d_segs.p, chunk, ncclDouble, g->comm, g->stream));
    if (g->coarse)
      if (gs_status s = coarse_gather(g, g->d_cvec.p, 3)) return s;
    GS_CUDA(g, cudaStreamSynchronize(g->stream));
  }
  const double t_comm = now_ms();
  g->prep_ms[2] = t_comm - t_up;
  if (gs_status s = capture_pcg_graph(g)) return s;
  g->prep_ms[3] = now_ms() - t_comm;
  if (ms_out) *ms_out = now_ms() - t0;
  return GS_OK;
}

// several ranks: vec holds `width` doubles per global aggregate; every rank broadcasts the share of its own aggregates
gs_status coarse_gather(gs_graph* g, double* vec, int width) {
  if (g->world == 1) return GS_OK;
  const gs::Aggregation& A = g->agg;
  const int spr = gs::kSegments / g->world;
  gs::NcclApi& nc = gs::nccl();
  GS_NCCL(g, nc.GroupStart());
  for (int q = 0; q < g->world; ++q) {
    const int b = A.seg_agg[q * spr], e = A.seg_agg[(q + 1) * spr];
    if (e > b)
      GS_NCCL(g, nc.Broadcast(vec + (size_t)b * width, vec + (size_t)b * width, (size_t)(e - b) * width, ncclDouble, q, g->comm,
                              g->stream));
  }
  GS_NCCL(g, nc.GroupEnd());
  return GS_OK;
}

// Two-level preconditioner, setup for the solve with damping `lambda` at the current linearisation:
//   E0 = W^T H W by columns — one undamped product H w per (colour, rigid mode), restricted per aggregate — and
//   G = W^T W, once per linearisation;  E = E0 + lambda G, Cholesky-factored and inverted (coarse_inverse: a dense
//   3 n_agg square, 3072 on C3, at most 12,288), once per lambda.  d_E then holds E^-1 in full.
gs_status coarse_setup(gs_graph* g, double lambda) {
  NvtxRange nvtx_range("gs:coarse setup");
  const double t0 = now_ms();
  const gs::Aggregation& A = g->agg;
  const int m = 3 * A.n_agg;
  if (!g->coarse_e0_valid) {
    GS_CUDA(g, cudaMemcpyAsync(g->d_pose_lin.p, g->d_pose.p, sizeof(double4) * (size_t)g->n_loc, cudaMemcpyDeviceToDevice, g->stream));
    const gs::CoarseView cv = coarse_view(g);
    const int nc9 = A.n_colors * 9;
    const int own = std::max(g->agg_own, 1);
    gs::k_agg_ref<<<(own + gs::kThreads - 1) / gs::kThreads, gs::kThreads, 0, g->stream>>>(cv, g->d_agg_ref.p);
    // several ranks: every rank learns the reference position of every aggregate once; the products below then form the
    // blocks of W of ghost columns themselves (from the ghost poses the rank holds) — no exchange per product
    if (gs_status s = coarse_gather(g, reinterpret_cast<double*>(g->d_agg_ref.p), 2)) return s;
    for (int c = 0; c < A.n_colors; ++c) {
      gs::k_coarse_product<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(spmv_view(g), cv, g->d_agg_color.p, c, g->d_Y.p);
      if (g->agg_own) gs::k_restrict_modes<<<g->agg_own, gs::kThreads, 0, g->stream>>>(cv, g->d_Y.p, g->d_R.p, nc9, c * 9);
    }
    for (int mode = 0; mode < 3; ++mode) {
      gs::k_basis_vec<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(cv, g->n_own, g->d_agg_color.p, -1, mode, g->d_basis.p);
      if (g->agg_own) gs::k_restrict<<<g->agg_own, gs::kThreads, 0, g->stream>>>(cv, g->d_basis.p, g->d_sc.p, 0, g->d_G.p, 9, mode * 3);
    }
    // every rank needs the whole coarse matrix: its aggregates' rows of R and G are contiguous (aggregate-major)
    if (gs_status s = coarse_gather(g, g->d_R.p, nc9)) return s;
    if (gs_status s = coarse_gather(g, g->d_G.p, 9)) return s;
    g->launches += 2 * (int64_t)A.n_colors + 7;
    g->coarse_e0_valid = true;
    g->coarse_lambda = -1.0;
  }
  if (g->coarse_lambda != lambda) {
    GS_CUDA(g, cudaMemsetAsync(g->d_E.p, 0, sizeof(double) * (size_t)m * m, g->stream));
    const int total = A.n_agg * A.n_colors * 3;
    gs::k_coarse_E<<<(total + gs::kThreads - 1) / gs::kThreads, gs::kThreads, 0, g->stream>>>(
        g->d_R.p, g->d_G.p, g->d_nbr_of_color.p, A.n_agg, A.n_colors, lambda, g->d_E.p);
    std::lock_guard<std::mutex> lock(solver_pool().mu);
    int info = 0;
    if (gs_status s = coarse_inverse(g, m, g->d_E.p, g->d_Ework.p, &info)) return s;
    if (info != 0)
      return g->fail(GS_ENAN, "coarse matrix of the two-level preconditioner is not positive definite (leading minor %d): is the "
                              "graph anchored by a prior?", info);
    g->coarse_lambda = lambda;
  }
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  GS_CUDA(g, cudaGetLastError());
  g->ms_coarse_setup += now_ms() - t0;
  return GS_OK;
}

gs_status read_scalars(gs_graph* g) {
  GS_CUDA(g, cudaMemcpyAsync(g->h_sc, g->d_sc.p, sizeof(Scalars), cudaMemcpyDeviceToHost, g->stream));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  g->d2h_bytes += sizeof(Scalars);
  return GS_OK;
}

gs_status write_scalars(gs_graph* g) {
  GS_CUDA(g, cudaMemcpyAsync(g->d_sc.p, g->h_sc, sizeof(Scalars), cudaMemcpyHostToDevice, g->stream));
  g->h2d_bytes += sizeof(Scalars);
  return GS_OK;
}

// linearise + assemble at the current device poses (async)
gs_status launch_assemble(gs_graph* g) {
  NvtxRange nvtx_range("gs:linearise+assemble");
  const gs::AsmView av = asm_view(g, g->d_pose.p);
  const gs::ReduceCtx rc = reduce_ctx(g, 2);
  // no vertex pair carries two factors (every BASELINE config): the form without the run accumulator, 80 registers and
  // three CTAs per SM instead of 118 and two (C3: 175.9 -> 163.5 us; same bits, tests/test_emulator_cpu.py)
  const bool nodup = g->S.n_merged == 0;
  if (g->prm.jacobian_mode == GS_JACOBIAN_EXACT) {
    if (nodup) gs::k_assemble_nodup<1><<<g->grid_rows, gs::kThreads, 0, g->stream>>>(av, rc);
    else gs::k_assemble<1><<<g->grid_rows, gs::kThreads, 0, g->stream>>>(av, rc);
  } else {
    if (nodup) gs::k_assemble_nodup<0><<<g->grid_rows, gs::kThreads, 0, g->stream>>>(av, rc);
    else gs::k_assemble<0><<<g->grid_rows, gs::kThreads, 0, g->stream>>>(av, rc);
  }
  g->launches += 1;
  if (gs_status s = finish_reduce(g, 1, gs::OP_CHI2)) return s;
  if (g->prm.diagonal_damping) {
    gs::k_dscale<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(g->d_Hd.p, g->d_dscale.p, g->n_own);
    g->launches += 1;
  }
  GS_CUDA(g, cudaGetLastError());
  g->system_valid = true;
  g->coarse_e0_valid = false;
  g->coarse_lambda = -1.0;
  return GS_OK;
}

// chi2 at `pose` (ghost slots must be current)
gs_status launch_chi2(gs_graph* g, const double4* pose) {
  NvtxRange nvtx_range("gs:chi2");
  // rows.cuh chi2_row: how far the slot loop runs ahead; GS_CHI2_PIPE (a compile-time switch of the variant probe) forces one
#ifndef GS_CHI2_PIPE
#define GS_CHI2_PIPE -1
#endif
  const int pipe = GS_CHI2_PIPE >= 0 ? GS_CHI2_PIPE : ((long long)g->e_count >= 8ll * g->n_own ? 2 : 1);
  if (pipe == 2) gs::k_chi2<2><<<g->grid_rows, gs::kThreads, 0, g->stream>>>(asm_view(g, pose), reduce_ctx(g, 3));
  else if (pipe == 1) gs::k_chi2<1><<<g->grid_rows, gs::kThreads, 0, g->stream>>>(asm_view(g, pose), reduce_ctx(g, 3));
  else gs::k_chi2<0><<<g->grid_rows, gs::kThreads, 0, g->stream>>>(asm_view(g, pose), reduce_ctx(g, 3));
  g->launches += 1;
  if (gs_status s = finish_reduce(g, 1, gs::OP_CHI2)) return s;
  GS_CUDA(g, cudaGetLastError());
  return GS_OK;
}

gs_status launch_pcg_setup(gs_graph* g) {
  gs::k_minv<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(g->d_Hd.p, g->prm.diagonal_damping ? g->d_dscale.p : nullptr,
                                                           g->d_sc.p, g->d_Minv.p, g->n_own);
  const bool two = coarse_active(g);
  gs::k_pcg_init<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(pcg_view(g), reduce_ctx(g, 1), two ? gs::OP_INIT_2L : gs::OP_INIT);
  g->launches += 2;
  if (two) {   // p(0) = z(0) = M^-1 r + W E^-1 W^T r
    const int m = 3 * g->agg.n_agg;
    if (gs_status s = finish_reduce(g, 2, gs::OP_INIT_2L)) return s;
    if (g->agg_own)
      gs::k_restrict<<<g->agg_own, gs::kThreads, 0, g->stream>>>(coarse_view(g), g->d_r.p, g->d_sc.p, 1, g->d_cvec.p, 3, 0);
    if (gs_status s = coarse_gather(g, g->d_cvec.p, 3)) return s;
    gs::k_coarse_mv<<<(m + gs::kCoarseRows - 1) / gs::kCoarseRows, gs::kThreads, 0, g->stream>>>(
        g->d_E.p, g->d_cvec.p, m, g->d_mu.p, g->d_cpart.p, g->d_coarse_seg_row.p, g->d_counter.p + 6, g->d_sc.p, 1);
    gs::k_dir_coarse<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(coarse_view(g), g->d_z.p, g->d_mu.p, g->pvec, g->d_sc.p);
    g->launches += 3;
    return GS_OK;
  }
  return finish_reduce(g, 2, gs::OP_INIT);
}

// peer-memory path: ghost rows of p(0) by the ordinary halo exchange, then mark them valid
gs_status begin_p2p_solve(gs_graph* g) {
  if (!g->p2p) return GS_OK;
  if (g->dir_otf) {
    // the window's ghost rows carry z here (p = z at the start); both ghost copies of p must be finite
    const size_t own3 = 3 * (size_t)g->n_own, gh = sizeof(double) * 3 * (size_t)std::max(g->S.n_ghost, 1);
    GS_CUDA(g, cudaMemsetAsync(g->pvec + own3, 0, gh, g->stream));
    GS_CUDA(g, cudaMemsetAsync(g->d_p_alt.p + own3, 0, gh, g->stream));
    if (gs_status s = halo_exchange(g, g->d_z.p, 3, g->pghost)) return s;
  } else if (gs_status s = halo_exchange(g, g->pvec, 3, g->pghost)) {
    return s;
  }
  gs::k_p2p_begin<<<1, 32, 0, g->stream>>>(reinterpret_cast<gs::P2PWindow*>(g->win), g->d_seq.p, g->d_sc.p);
  g->launches += 1;
  return GS_OK;
}

int64_t launches_per_pcg_iteration(const gs_graph* g) {
  if (g->dir_otf && g->p2p) return 2;
  if (coarse_active(g) && g->p2p) return 6;
  if (coarse_active(g)) return g->world == 1 ? 5 : 5 + 2 + (g->S.send_rows.empty() ? 0 : 1);
  if (g->world == 1 || g->p2p) return 3;
  return 3 + 2 + (g->S.send_rows.empty() ? 0 : 1);  // + two k_scalars + k_pack
}

// Solve (H + lambda D) x = -b by block-Jacobi PCG; result in d_x.  H, b must be assembled.
gs_status pcg_solve(gs_graph* g, double lambda, int* iters_out, double* relres_out, int* done_out) {
  NvtxRange nvtx_range("gs:pcg solve");
  Scalars* h = g->h_sc;
  if (coarse_active(g))
    if (gs_status s = coarse_setup(g, lambda)) return s;
  std::memset(h, 0, sizeof(Scalars));
  h->lambda = lambda;
  h->tol2 = g->prm.pcg_rel_tol * g->prm.pcg_rel_tol;
  h->max_iters = g->prm.pcg_max_iters;
  if (gs_status s = write_scalars(g)) return s;
  if (gs_status s = launch_pcg_setup(g)) return s;
  if (gs_status s = begin_p2p_solve(g)) return s;
  GS_CUDA(g, cudaGetLastError());
  const int64_t keep = g->launches;
  int64_t fused_launches = 0;
  for (;;) {
    if (g->fused) {
      if (gs_status s = launch_pcg_fused(g, kFusedBatch)) return s;
      ++fused_launches;
    } else if (g->pcg_exec) {
      GS_CUDA(g, cudaGraphLaunch(g->pcg_exec, g->stream));
    } else {
      if (gs_status s = launch_pcg_batch(g, kPcgBatch)) return s;
    }
    if (gs_status s = read_scalars(g)) return s;
    if (g->comm) {   // a failed peer / transport error surfaces here instead of as a hang in the next batch
      ncclResult_t async = ncclSuccess;
      if (gs::nccl().CommGetAsyncError(g->comm, &async) != ncclSuccess || async != ncclSuccess)
        return g->fail(GS_ENCCL, "NCCL asynchronous error during PCG: %s", gs::nccl().GetErrorString(async));
    }
    if (h->done) break;
  }
  if (g->p2p)
    for (int k = 0; k < 4; ++k) g->p2p_wait[k] += (double)h->spin[k];
  if (h->done == 4) return g->fail(GS_ENCCL, "a peer rank did not answer within the wait limit (peer-memory PCG)");
  // kernels that did work (the no-op tail of the last batch is not counted)
  g->launches = keep + (g->fused ? fused_launches : launches_per_pcg_iteration(g) * (int64_t)h->iters);
  if (iters_out) *iters_out = h->iters;
  if (relres_out) *relres_out = h->rr0 > 0.0 ? std::sqrt(h->rr / h->rr0) : 0.0;
  if (done_out) *done_out = h->done;
  return GS_OK;
}

float elapsed(gs_graph* g) {
  float ms = 0.f;
  cudaEventElapsedTime(&ms, g->ev0, g->ev1);
  return ms;
}

gs_status create_common(gs_graph* g) {
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return g->fail(GS_ECUDA, "no CUDA device (%s); libgraphslam_cuda has no CPU fallback",
                   e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (g->prm.device >= 0) GS_CUDA(g, cudaSetDevice(g->prm.device));
  GS_CUDA(g, cudaStreamCreateWithFlags(&g->stream, cudaStreamNonBlocking));
  GS_CUDA(g, cudaEventCreate(&g->ev0));
  GS_CUDA(g, cudaEventCreate(&g->ev1));
  GS_CUDA(g, cudaEventCreate(&g->ev_begin));
  GS_CUDA(g, cudaEventCreate(&g->ev_end));
  GS_CUDA(g, g->d_sc.alloc(1));
  GS_CUDA(g, g->d_counter.alloc(8));
  GS_CUDA(g, g->d_segs.alloc(2 * gs::kSegments));
  GS_CUDA(g, cudaMemsetAsync(g->d_counter.p, 0, 8 * sizeof(unsigned int), g->stream));
  GS_CUDA(g, cudaMemsetAsync(g->d_sc.p, 0, sizeof(Scalars), g->stream));
  GS_CUDA(g, cudaMemsetAsync(g->d_segs.p, 0, 2 * gs::kSegments * sizeof(double), g->stream));
  GS_CUDA(g, cudaMallocHost(&g->h_sc, sizeof(Scalars)));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  const int flags = g->prm.solver_flags;
  g->use_graph = !(flags & GS_SOLVER_NO_CUDA_GRAPH);
  g->fused_wanted = !(flags & GS_SOLVER_NO_PERSISTENT);
  g->p2p_wanted = !(flags & GS_SOLVER_NCCL_PER_OP);
  g->coarse_wanted = !(flags & GS_SOLVER_BLOCK_JACOBI_ONLY);
  {
    int per_sm = 0, sms = 0, dev = 0, coop = 0;
    GS_CUDA(g, cudaGetDevice(&dev));
    GS_CUDA(g, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    GS_CUDA(g, cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev));
    GS_CUDA(g, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, gs::k_pcg_fused, gs::kThreads, 0));
    g->fused_capacity = coop ? per_sm * sms : 0;
  }
  GS_CUDA(g, g->d_bar.alloc(1));
  GS_CUDA(g, cudaMemsetAsync(g->d_bar.p, 0, sizeof(gs::GridBarrier), g->stream));
  {
    // a wait for another GPU gives up after peer_wait_limit_ms (default 30 s: long enough for a descheduled or
    // page-faulting peer process, short enough to turn a dead peer into GS_ENCCL instead of a hang)
    int dev = 0, khz = 0;
    GS_CUDA(g, cudaGetDevice(&dev));
    GS_CUDA(g, cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev));
    const long long ms = g->prm.peer_wait_limit_ms > 0 ? g->prm.peer_wait_limit_ms : 30000;
    g->spin_limit_ticks = ms * (long long)std::max(khz, 1000000);
  }
  GS_CUDA(g, g->d_seq.alloc(64));
  GS_CUDA(g, cudaMemsetAsync(g->d_seq.p, 0, 64 * sizeof(unsigned int), g->stream));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  return GS_OK;
}

}  // namespace

// No exception crosses the C ABI: every entry point is a function-try-block that ends here.  The host side
// allocates (std::vector growth while adding factors, the structure build); std::bad_alloc and anything else a
// standard container can throw become a status, with the message stored when that itself does not throw.
static gs_status on_exception(gs_graph* g) noexcept {
  gs_status code = GS_EINVAL;
  const char* what = "unexpected C++ exception";
  try {
    throw;
  } catch (const std::bad_alloc&) {
    code = GS_ENOMEM;
    what = "out of host memory";
  } catch (const std::exception& e) {
    what = e.what();
    try {
      if (g) g->fail(code, "C++ exception: %s", what);
    } catch (...) {
    }
    return code;
  } catch (...) {
  }
  try {
    if (g) g->fail(code, "%s", what);
  } catch (...) {
  }
  return code;
}

// =================================================================================================
extern "C" {

void gs_default_params(gs_params* p) {
  if (!p) return;
  std::memset(p, 0, sizeof *p);
  p->max_iterations = 100;
  p->rel_error_tol = 1e-5;
  p->abs_error_tol = 1e-5;
  p->error_tol = 0.0;
  p->lambda_initial = 1e-5;
  p->lambda_factor = 10.0;
  p->lambda_upper = 1e5;
  p->lambda_lower = 0.0;
  p->min_model_fidelity = 1e-3;
  p->diagonal_damping = 0;
  p->jacobian_mode = GS_JACOBIAN_GTSAM_DEFAULT;
  p->pcg_rel_tol = 1e-12;
  p->pcg_max_iters = 100000;
  p->device = -1;
  p->rank = 0;
  p->world_size = 1;
}

int gs_version(void) { return GS_VERSION_MAJOR * 1000 + GS_VERSION_MINOR; }

gs_status gs_create(gs_graph** out, const gs_params* params) try {
  if (!out) return GS_EINVAL;
  *out = nullptr;
  gs_graph* g = new (std::nothrow) gs_graph();
  if (!g) return GS_EINVAL;
  if (params) g->prm = *params;
  else gs_default_params(&g->prm);
  *out = g;  // returned even on failure so that gs_last_error can be read; caller destroys it
  if (g->prm.world_size != 1 || g->prm.rank != 0)
    return g->fail(GS_EINVAL, "multi-rank graphs are created with gs_create_distributed");
  return create_common(g);
} catch (...) {
  return on_exception(nullptr);
}

gs_status gs_nccl_unique_id(void* out, size_t cap, size_t* n) try {
  if (n) *n = sizeof(ncclUniqueId);
  if (!out) return GS_OK;
  if (cap < sizeof(ncclUniqueId)) return GS_EINVAL;
  if (!gs::nccl().load().empty()) return GS_ENCCL;
  ncclUniqueId id;
  if (gs::nccl().GetUniqueId(&id) != ncclSuccess) return GS_ENCCL;
  std::memcpy(out, &id, sizeof id);
  return GS_OK;
} catch (...) {
  return on_exception(nullptr);
}

gs_status gs_create_distributed(gs_graph** out, const gs_params* params, const void* nccl_id, size_t id_bytes) try {
  if (!out) return GS_EINVAL;
  *out = nullptr;
  gs_graph* g = new (std::nothrow) gs_graph();
  if (!g) return GS_EINVAL;
  if (params) g->prm = *params;
  else gs_default_params(&g->prm);
  *out = g;
  const int w = g->prm.world_size, r = g->prm.rank;
  if (!(w == 1 || w == 2 || w == 4 || w == 8) || r < 0 || r >= w)
    return g->fail(GS_EINVAL, "world_size must be 1, 2, 4 or 8 and 0 <= rank < world_size (got rank %d of %d)", r, w);
  g->world = w;
  g->rank = r;
  if (gs_status s = create_common(g)) return s;
  if (w > 1) {
    if (!nccl_id || id_bytes != sizeof(ncclUniqueId)) return g->fail(GS_EINVAL, "nccl_id must be %zu bytes", sizeof(ncclUniqueId));
    const std::string why = gs::nccl().load();
    if (!why.empty()) return g->fail(GS_ENCCL, "%s", why.c_str());
    ncclUniqueId id;
    std::memcpy(&id, nccl_id, sizeof id);
    GS_NCCL(g, gs::nccl().CommInitRank(&g->comm, w, id, r));
  }
  return GS_OK;
} catch (...) {
  return on_exception(nullptr);
}

gs_status gs_destroy(gs_graph* g) try {
  if (!g) return GS_EINVAL;
  if (g->stream) cudaStreamSynchronize(g->stream);
  destroy_graph_exec(g);
  for (int q = 0; q < gs::kMaxRanks; ++q) close_peer(g, q);
  if (g->comm && g->win) {
    // nobody frees a window a peer may still have mapped: one last all-gather as a barrier (gs_destroy is collective)
    XchgRecord mine;
    std::memset(&mine, 0, sizeof mine);
    std::vector<XchgRecord> all;
    exchange_records(g, mine, all);
  }
  if (g->win) cudaFree(g->win);
  if (g->comm) gs::nccl().CommDestroy(g->comm);
  if (g->h_sc) cudaFreeHost(g->h_sc);
  if (g->h_stage) cudaFreeHost(g->h_stage);
  for (int b = 0; b < 2; ++b) {
    if (g->h_struct[b]) cudaFreeHost(g->h_struct[b]);
    if (g->ev_struct[b]) cudaEventDestroy(g->ev_struct[b]);
  }
  if (g->ev0) cudaEventDestroy(g->ev0);
  if (g->ev1) cudaEventDestroy(g->ev1);
  if (g->ev_begin) cudaEventDestroy(g->ev_begin);
  if (g->ev_end) cudaEventDestroy(g->ev_end);
  if (g->stream) cudaStreamDestroy(g->stream);
  delete g;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_set_params(gs_graph* g, const gs_params* p) try {
  if (!g || !p) return GS_EINVAL;
  if (p->world_size != g->prm.world_size || p->rank != g->prm.rank || p->device != g->prm.device ||
      p->solver_flags != g->prm.solver_flags)
    return g->fail(GS_EINVAL, "device/rank/world_size/solver_flags are fixed at gs_create");
  const bool mode_changed = p->jacobian_mode != g->prm.jacobian_mode || p->diagonal_damping != g->prm.diagonal_damping;
  const bool damping_changed = p->diagonal_damping != g->prm.diagonal_damping;
  g->prm = *p;
  if (mode_changed) g->system_valid = false;
  // the product's dscale pointer (lambda*I or lambda*diag(H)) is an argument captured in the PCG graph
  if (damping_changed && g->pcg_exec)
    if (gs_status s = capture_pcg_graph(g)) return s;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_add_vertex(gs_graph* g, uint64_t key, const double pose[3]) try {
  if (!g || !pose) return GS_EINVAL;
  if (!(std::isfinite(pose[0]) && std::isfinite(pose[1]) && std::isfinite(pose[2])))
    return g->fail(GS_EINVAL, "non-finite pose for key %llu", (unsigned long long)key);
  if (g->find(key) >= 0) return g->fail(GS_EDUPKEY, "key %llu already present", (unsigned long long)key);
  if (gs_status s = download_values(g)) return s;
  const int idx = (int)g->keys.size();
  if (g->keys_sequential && idx > 0 && key != g->keys[0] + (uint64_t)idx) {
    g->keys_sequential = false;
    g->key2ins.reserve(g->keys.size() * 2);
    for (int k = 0; k < idx; ++k) g->key2ins[g->keys[k]] = k;
  }
  if (!g->keys_sequential) g->key2ins[key] = idx;
  g->keys.push_back(key);
  g->theta_wrapped = false;
  g->xyt.insert(g->xyt.end(), pose, pose + 3);
  g->topo_dirty = true;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_add_prior(gs_graph* g, uint64_t key, const double pose[3], const double cov9[9]) try {
  if (!g || !pose || !cov9) return GS_EINVAL;
  const int v = g->find(key);
  if (v < 0) return g->fail(GS_ENOKEY, "prior on unknown key %llu", (unsigned long long)key);
  if (!(std::isfinite(pose[0]) && std::isfinite(pose[1]) && std::isfinite(pose[2]))) return g->fail(GS_EINVAL, "non-finite prior");
  double om[6];
  if (!spd_information(cov9, om)) return g->fail(GS_ENOTSPD, "prior covariance on key %llu is not SPD", (unsigned long long)key);
  g->pr_v.push_back(v);
  g->pr_xyt.insert(g->pr_xyt.end(), pose, pose + 3);
  g->pr_om.insert(g->pr_om.end(), om, om + 6);
  g->topo_dirty = true;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_add_between(gs_graph* g, uint64_t k1, uint64_t k2, const double d[3], const double cov9[9]) try {
  if (!g || !d || !cov9) return GS_EINVAL;
  const int a = g->find(k1), b = g->find(k2);
  if (a < 0 || b < 0)
    return g->fail(GS_ENOKEY, "between factor names unknown key %llu", (unsigned long long)(a < 0 ? k1 : k2));
  if (a == b) return g->fail(GS_EINVAL, "between factor from key %llu to itself", (unsigned long long)k1);
  if (!(std::isfinite(d[0]) && std::isfinite(d[1]) && std::isfinite(d[2]))) return g->fail(GS_EINVAL, "non-finite delta");
  double om[6];
  if (!spd_information(cov9, om))
    return g->fail(GS_ENOTSPD, "covariance of factor %llu->%llu is not SPD", (unsigned long long)k1, (unsigned long long)k2);
  g->e1.push_back(a);
  g->e2.push_back(b);
  g->ez.insert(g->ez.end(), d, d + 3);
  g->eom.insert(g->eom.end(), om, om + 6);
  g->topo_dirty = true;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

// noiseModel::Gaussian::Information(I) forms of the two calls above (what a g2o file carries)
gs_status gs_add_prior_information(gs_graph* g, uint64_t key, const double pose[3], const double info6[6]) try {
  if (!g || !pose || !info6) return GS_EINVAL;
  const int v = g->find(key);
  if (v < 0) return g->fail(GS_ENOKEY, "prior on unknown key %llu", (unsigned long long)key);
  if (!(std::isfinite(pose[0]) && std::isfinite(pose[1]) && std::isfinite(pose[2]))) return g->fail(GS_EINVAL, "non-finite prior");
  if (!spd_upper(info6)) return g->fail(GS_ENOTSPD, "prior information on key %llu is not SPD", (unsigned long long)key);
  g->pr_v.push_back(v);
  g->pr_xyt.insert(g->pr_xyt.end(), pose, pose + 3);
  g->pr_om.insert(g->pr_om.end(), info6, info6 + 6);
  g->topo_dirty = true;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_add_between_information(gs_graph* g, uint64_t k1, uint64_t k2, const double d[3], const double info6[6]) try {
  if (!g || !d || !info6) return GS_EINVAL;
  const int a = g->find(k1), b = g->find(k2);
  if (a < 0 || b < 0)
    return g->fail(GS_ENOKEY, "between factor names unknown key %llu", (unsigned long long)(a < 0 ? k1 : k2));
  if (a == b) return g->fail(GS_EINVAL, "between factor from key %llu to itself", (unsigned long long)k1);
  if (!(std::isfinite(d[0]) && std::isfinite(d[1]) && std::isfinite(d[2]))) return g->fail(GS_EINVAL, "non-finite delta");
  if (!spd_upper(info6))
    return g->fail(GS_ENOTSPD, "information of factor %llu->%llu is not SPD", (unsigned long long)k1, (unsigned long long)k2);
  g->e1.push_back(a);
  g->e2.push_back(b);
  g->ez.insert(g->ez.end(), d, d + 3);
  g->eom.insert(g->eom.end(), info6, info6 + 6);
  g->topo_dirty = true;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

// g2o / TORO text (csrc/g2o.hpp).  Loading appends to the graph: vertices first, then priors, then
// edges, in file order; nothing is added when the file is malformed or names an unknown/duplicate key.
gs_status gs_load_g2o(gs_graph* g, const char* path, gs_g2o_counts* counts) try {
  if (!g || !path) return GS_EINVAL;
  gs::G2oData d;
  const std::string err = gs::read_g2o(path, d);
  if (!err.empty()) return g->fail(GS_EINVAL, "%s", err.c_str());
  // validate everything before touching the graph
  {
    std::unordered_map<uint64_t, int> seen;
    seen.reserve(d.vkey.size() * 2);
    for (uint64_t k : d.vkey) {
      if (g->find(k) >= 0 || !seen.emplace(k, 1).second)
        return g->fail(GS_EDUPKEY, "%s: vertex %llu already present", path, (unsigned long long)k);
    }
    auto known = [&](uint64_t k) { return seen.count(k) || g->find(k) >= 0; };
    for (size_t k = 0; k < d.e1.size(); ++k) {
      if (!known(d.e1[k]) || !known(d.e2[k]))
        return g->fail(GS_ENOKEY, "%s: edge %zu names an unknown vertex", path, k);
      if (d.e1[k] == d.e2[k]) return g->fail(GS_EINVAL, "%s: edge %zu joins a vertex to itself", path, k);
      if (!spd_upper(&d.einfo[6 * k])) return g->fail(GS_ENOTSPD, "%s: information of edge %zu is not SPD", path, k);
      if (!(std::isfinite(d.ez[3 * k]) && std::isfinite(d.ez[3 * k + 1]) && std::isfinite(d.ez[3 * k + 2])))
        return g->fail(GS_EINVAL, "%s: non-finite measurement in edge %zu", path, k);
    }
    for (size_t k = 0; k < d.pkey.size(); ++k) {
      if (!known(d.pkey[k])) return g->fail(GS_ENOKEY, "%s: prior %zu names an unknown vertex", path, k);
      if (!spd_upper(&d.pinfo[6 * k])) return g->fail(GS_ENOTSPD, "%s: information of prior %zu is not SPD", path, k);
      if (!(std::isfinite(d.pxyt[3 * k]) && std::isfinite(d.pxyt[3 * k + 1]) && std::isfinite(d.pxyt[3 * k + 2])))
        return g->fail(GS_EINVAL, "%s: non-finite pose in prior %zu", path, k);
    }
    for (size_t k = 0; k < d.vkey.size(); ++k)
      if (!(std::isfinite(d.vxyt[3 * k]) && std::isfinite(d.vxyt[3 * k + 1]) && std::isfinite(d.vxyt[3 * k + 2])))
        return g->fail(GS_EINVAL, "%s: non-finite pose for vertex %llu", path, (unsigned long long)d.vkey[k]);
  }
  g->keys.reserve(g->keys.size() + d.vkey.size());
  g->xyt.reserve(g->xyt.size() + d.vxyt.size());
  for (size_t k = 0; k < d.vkey.size(); ++k)
    if (gs_status s = gs_add_vertex(g, d.vkey[k], &d.vxyt[3 * k])) return s;
  for (size_t k = 0; k < d.pkey.size(); ++k)
    if (gs_status s = gs_add_prior_information(g, d.pkey[k], &d.pxyt[3 * k], &d.pinfo[6 * k])) return s;
  g->e1.reserve(g->e1.size() + d.e1.size());
  g->e2.reserve(g->e2.size() + d.e1.size());
  g->ez.reserve(g->ez.size() + d.ez.size());
  g->eom.reserve(g->eom.size() + d.einfo.size());
  for (size_t k = 0; k < d.e1.size(); ++k)
    if (gs_status s = gs_add_between_information(g, d.e1[k], d.e2[k], &d.ez[3 * k], &d.einfo[6 * k])) return s;
  if (counts) {
    counts->vertices = d.vkey.size();
    counts->edges = d.e1.size();
    counts->priors = d.pkey.size();
    counts->fixed = d.fixed.size();
    counts->skipped_lines = d.skipped;
  }
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

// the whole graph with the CURRENT values (after gs_optimize: the optimum), loadable by gs_load_g2o
gs_status gs_save_g2o(gs_graph* g, const char* path) try {
  if (!g || !path) return GS_EINVAL;
  if (gs_status s = download_values(g)) return s;
  gs::G2oData d;
  d.vkey = g->keys;
  d.vxyt = g->xyt;
  d.e1.resize(g->e1.size());
  d.e2.resize(g->e2.size());
  for (size_t k = 0; k < g->e1.size(); ++k) {
    d.e1[k] = g->keys[g->e1[k]];
    d.e2[k] = g->keys[g->e2[k]];
  }
  d.ez = g->ez;
  d.einfo = g->eom;
  d.pkey.resize(g->pr_v.size());
  for (size_t k = 0; k < g->pr_v.size(); ++k) d.pkey[k] = g->keys[g->pr_v[k]];
  d.pxyt = g->pr_xyt;
  d.pinfo = g->pr_om;
  const std::string err = gs::write_g2o(path, d);
  if (!err.empty()) return g->fail(GS_EINVAL, "%s", err.c_str());
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_add_batch(gs_graph* g, size_t nv, const uint64_t* keys, const double* poses, size_t ne, const uint64_t* k1,
                       const uint64_t* k2, const double* deltas, const double* cov9s) try {
  if (!g) return GS_EINVAL;
  if ((nv && (!keys || !poses)) || (ne && (!k1 || !k2 || !deltas || !cov9s))) return g->fail(GS_EINVAL, "null array");
  g->keys.reserve(g->keys.size() + nv);
  g->xyt.reserve(g->xyt.size() + 3 * nv);
  for (size_t k = 0; k < nv; ++k)
    if (gs_status s = gs_add_vertex(g, keys[k], poses + 3 * k)) return s;
  g->e1.reserve(g->e1.size() + ne);
  g->e2.reserve(g->e2.size() + ne);
  g->ez.reserve(g->ez.size() + 3 * ne);
  g->eom.reserve(g->eom.size() + 6 * ne);
  for (size_t k = 0; k < ne; ++k)
    if (gs_status s = gs_add_between(g, k1[k], k2[k], deltas + 3 * k, cov9s + 9 * k)) return s;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_prepare(gs_graph* g) try {
  if (!g) return GS_EINVAL;
  return prepare(g, nullptr);
} catch (...) {
  return on_exception(g);
}

gs_status gs_get_structure_timing(const gs_graph* g, double ms[4]) try {
  if (!g || !ms) return GS_EINVAL;
  for (int k = 0; k < 4; ++k) ms[k] = g->prep_ms[k];
  return GS_OK;
} catch (...) {
  return on_exception(const_cast<gs_graph*>(g));
}

gs_status gs_optimize(gs_graph* g, gs_stats* st) try {
  if (!g) return GS_EINVAL;
  const double t_begin = now_ms();
  gs_stats local;
  std::memset(&local, 0, sizeof local);
  double ms_structure = 0.0;
  const double h2d0 = g->h2d_bytes, d2h0 = g->d2h_bytes;
  if (gs_status s = prepare(g, &ms_structure)) return s;
  local.ms_structure = ms_structure;
  if (ms_structure > 0.0)
    for (int k = 0; k < 4; ++k) local.ms_structure_parts[k] = g->prep_ms[k];
  GS_CUDA(g, cudaEventRecord(g->ev_begin, g->stream));
  const gs_params& P = g->prm;
  Scalars* h = g->h_sc;
  g->tr_err.clear();
  g->tr_lambda.clear();
  g->tr_tries.clear();
  g->launches = 0;
  g->ms_coarse_setup = 0.0;
  for (double& w : g->p2p_wait) w = 0.0;

  // error at the initial values
  if (gs_status s = launch_chi2(g, g->d_pose.p)) return s;
  if (gs_status s = read_scalars(g)) return s;
  double error = 0.5 * h->chi2;
  if (!std::isfinite(error)) return g->fail(GS_ENAN, "initial chi2 is not finite");
  local.chi2_initial = 2.0 * error;
  g->tr_err.push_back(error);
  double lambda = P.lambda_initial;
  int iterations = 0;
  gs_status rc = GS_OK;
  const double eps = std::numeric_limits<double>::epsilon();

  // NonlinearOptimizer::defaultOptimize: nothing to do when the initial error is already below errorTol
  const bool nothing_to_do = error <= P.error_tol;
  if (nothing_to_do) local.converged = 1;
  while (!nothing_to_do && iterations < P.max_iterations) {
    const double current_error = error;
    // ---- iterate(): linearise once
    GS_CUDA(g, cudaEventRecord(g->ev0, g->stream));
    if (gs_status s = launch_assemble(g)) return s;
    GS_CUDA(g, cudaEventRecord(g->ev1, g->stream));
    GS_CUDA(g, cudaEventSynchronize(g->ev1));
    local.ms_linearize += elapsed(g);
    local.outer_loops += 1;
    int tries = 0;
    for (;;) {  // tryLambda
      ++tries;
      local.inner_iterations += 1;
      GS_CUDA(g, cudaEventRecord(g->ev0, g->stream));
      int pit = 0, done = 0;
      double relres = 0.0;
      if (gs_status s = pcg_solve(g, lambda, &pit, &relres, &done)) return s;
      GS_CUDA(g, cudaEventRecord(g->ev1, g->stream));
      GS_CUDA(g, cudaEventSynchronize(g->ev1));
      local.ms_solve += elapsed(g);
      local.pcg_iterations_total += pit;
      const bool solved = (done == 1);
      if (done == 3) rc = GS_ENOTCONVERGED;
      bool step_ok = false, stop_search = false;
      double new_error = std::numeric_limits<double>::infinity();
      if (solved || done == 3) {
        // linear.error(0) - linear.error(delta) = -delta.b - 0.5 delta.H.delta  (undamped H)
        GS_CUDA(g, cudaEventRecord(g->ev0, g->stream));
        if (gs_status s = launch_spmv(g, g->d_x.p, g->d_Ap.p, 1)) return s;
        if (gs_status s = read_scalars(g)) return s;
        const double dHd = h->dot;
        gs::k_dot<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(g->d_x.p, g->d_b.p, g->n_own, reduce_ctx(g, 4));
        if (gs_status s = finish_reduce(g, 1, gs::OP_DOT)) return s;
        if (gs_status s = read_scalars(g)) return s;
        const double db = h->dot;
        g->launches += 2;
        const double lin_change = -db - 0.5 * dHd;
        if (lin_change >= 0.0) {
          gs::k_retract<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(g->d_pose.p, g->d_x.p, g->d_pose_try.p, g->n_own);
          g->launches += 1;
          if (gs_status s = halo_exchange(g, reinterpret_cast<double*>(g->d_pose_try.p), 4)) return s;
          if (gs_status s = launch_chi2(g, g->d_pose_try.p)) return s;
          if (gs_status s = read_scalars(g)) return s;
          new_error = 0.5 * h->chi2;
          const double cost_change = error - new_error;
          if (lin_change > eps * error) {
            const double fidelity = cost_change / lin_change;
            step_ok = fidelity > P.min_model_fidelity;
          }
          if (std::fabs(cost_change) < P.rel_error_tol * error) stop_search = true;
        }
        GS_CUDA(g, cudaEventRecord(g->ev1, g->stream));
        GS_CUDA(g, cudaEventSynchronize(g->ev1));
        local.ms_update += elapsed(g);
      }
      if (step_ok) {
        g->tr_lambda.push_back(lambda);
        std::swap(g->d_pose.p, g->d_pose_try.p);  // accept: candidate becomes current
        std::swap(g->d_pose.n, g->d_pose_try.n);
        g->host_values_fresh = false;
        g->system_valid = false;
        error = new_error;
        lambda = std::max(P.lambda_lower, lambda / P.lambda_factor);
        ++iterations;
        break;
      } else if (!stop_search) {
        lambda *= P.lambda_factor;
        if (lambda >= P.lambda_upper) break;
      } else {
        break;
      }
    }
    g->tr_tries.push_back(tries);
    g->tr_err.push_back(error);
    if (!std::isfinite(error)) {
      rc = g->fail(GS_ENAN, "chi2 became non-finite");
      break;
    }
    // ---- checkConvergence(relTol, absTol, errTol, currentError, newError)
    if (iterations >= P.max_iterations) break;
    if (error <= P.error_tol) { local.converged = 1; break; }
    const double abs_dec = current_error - error;
    const double rel_dec = current_error != 0.0 ? abs_dec / current_error : 0.0;
    if ((P.rel_error_tol != 0.0 && rel_dec <= P.rel_error_tol) || abs_dec <= P.abs_error_tol) { local.converged = 1; break; }
  }
  local.iterations = iterations;
  local.chi2_final = 2.0 * error;
  local.lambda_final = lambda;
  local.ms_total = now_ms() - t_begin;
  local.edges_per_sec = local.ms_linearize > 0 ? (double)g->E * local.outer_loops / (local.ms_linearize * 1e-3) : 0.0;
  const double t_iter = local.ms_total - local.ms_structure;
  local.gn_iters_per_sec = t_iter > 0 ? local.outer_loops / (t_iter * 1e-3) : 0.0;
  local.gpu_launches = (double)g->launches;
  local.ms_coarse_setup = g->ms_coarse_setup;
  GS_CUDA(g, cudaEventRecord(g->ev_end, g->stream));
  GS_CUDA(g, cudaEventSynchronize(g->ev_end));
  {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, g->ev_begin, g->ev_end);
    local.ms_device = ms;
  }
  local.h2d_bytes = g->h2d_bytes - h2d0;
  local.d2h_bytes = g->d2h_bytes - d2h0;
  for (int k = 0; k < 3; ++k) local.p2p_wait_ticks[k] = g->p2p_wait[k];
  local.p2p_waits = g->p2p_wait[3];
  if (st) *st = local;
  return rc;
} catch (...) {
  return on_exception(g);
}

gs_status gs_get_pose(gs_graph* g, uint64_t key, double out[3]) try {
  if (!g || !out) return GS_EINVAL;
  const int v = g->find(key);
  if (v < 0) return g->fail(GS_ENOKEY, "unknown key %llu", (unsigned long long)key);
  if (gs_status s = download_values(g)) return s;
  out[0] = g->xyt[3 * v];
  out[1] = g->xyt[3 * v + 1];
  const double th = g->xyt[3 * v + 2];
  out[2] = (th > -M_PI && th <= M_PI) ? th : std::atan2(std::sin(th), std::cos(th));
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_get_poses(gs_graph* g, size_t cap, uint64_t* keys, double* xyt, size_t* n) try {
  if (!g) return GS_EINVAL;
  const size_t nv = g->keys.size();
  if (n) *n = nv;
  if (!keys && !xyt) return GS_OK;
  if (cap < nv) return g->fail(GS_EINVAL, "capacity %zu < %zu vertices", cap, nv);
  if (gs_status s = download_values(g)) return s;
  if (g->dense2ins.size() != nv) build_dense_order(g);
  const bool wrapped = g->theta_wrapped;
  gs::parallel_for(nv, [g, keys, xyt, wrapped](size_t b, size_t e) {
    for (size_t d = b; d < e; ++d) {
      const int k = g->dense2ins[d];
      if (keys) keys[d] = g->keys[k];
      if (xyt) {
        const double th = g->xyt[3 * k + 2];
        xyt[3 * d] = g->xyt[3 * k];
        xyt[3 * d + 1] = g->xyt[3 * k + 1];
        // theta in (-pi, pi]: values already there are returned as they are (bit-exact save -> load -> read), only
        // others are wrapped (values read back from the device are atan2 results, always in range)
        xyt[3 * d + 2] = (wrapped || (th > -M_PI && th <= M_PI)) ? th : std::atan2(std::sin(th), std::cos(th));
      }
    }
  });
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

// ---- rank-local pose transfers: what a rank owns, and nothing else, crosses its PCIe link ------------------------
gs_status gs_local_range(gs_graph* g, size_t* first, size_t* count) try {
  if (!g) return GS_EINVAL;
  if (g->topo_dirty) return g->fail(GS_EINVAL, "gs_local_range: the structure is not built (call gs_prepare after the last gs_add_*)");
  if (first) *first = (size_t)g->S.v0;
  if (count) *count = (size_t)g->n_own;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_get_local_poses(gs_graph* g, size_t cap, uint64_t* keys, double* xyt, size_t* count) try {
  if (!g) return GS_EINVAL;
  if (g->topo_dirty) return g->fail(GS_EINVAL, "gs_get_local_poses: the structure is not built (call gs_prepare after the last gs_add_*)");
  const size_t n = (size_t)g->n_own, v0 = (size_t)g->S.v0;
  if (count) *count = n;
  if (!keys && !xyt) return GS_OK;
  if (cap < n) return g->fail(GS_EINVAL, "capacity %zu < %zu owned vertices", cap, n);
  if (keys) {
    gs::parallel_for(n, [g, keys, v0](size_t b, size_t e) {
      for (size_t l = b; l < e; ++l) keys[l] = g->keys[g->dense2ins[v0 + l]];
    });
  }
  if (!xyt) return GS_OK;
  if (g->device_values_valid && !g->host_values_fresh) {
    // the optimised values live on the device: this rank's rows only, through the pinned stage
    if (gs_status s = ensure_stage(g, std::max<size_t>(n, 1))) return s;
    double4* tmp = g->h_stage;
    if (n) GS_CUDA(g, cudaMemcpyAsync(tmp, g->d_pose.p, sizeof(double4) * n, cudaMemcpyDeviceToHost, g->stream));
    GS_CUDA(g, cudaStreamSynchronize(g->stream));
    g->d2h_bytes += sizeof(double4) * (double)n;
    gs::parallel_for(n, [tmp, xyt](size_t b, size_t e) {
      for (size_t l = b; l < e; ++l) {
        xyt[3 * l] = tmp[l].x;
        xyt[3 * l + 1] = tmp[l].y;
        xyt[3 * l + 2] = std::atan2(tmp[l].w, tmp[l].z);
      }
    });
  } else {
    const bool wrapped = g->theta_wrapped;
    gs::parallel_for(n, [g, xyt, v0, wrapped](size_t b, size_t e) {
      for (size_t l = b; l < e; ++l) {
        const int k = g->dense2ins[v0 + l];
        const double th = g->xyt[3 * k + 2];
        xyt[3 * l] = g->xyt[3 * k];
        xyt[3 * l + 1] = g->xyt[3 * k + 1];
        xyt[3 * l + 2] = (wrapped || (th > -M_PI && th <= M_PI)) ? th : std::atan2(std::sin(th), std::cos(th));
      }
    });
  }
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_set_local_poses(gs_graph* g, size_t n, const double* xyt) try {
  if (!g || (n && !xyt)) return GS_EINVAL;
  if (g->topo_dirty) return g->fail(GS_EINVAL, "gs_set_local_poses: the structure is not built (call gs_prepare after the last gs_add_*)");
  if (n != (size_t)g->n_own) return g->fail(GS_EINVAL, "gs_set_local_poses: %zu poses given, this rank owns %d", n, g->n_own);
  {
    // the other rows of the device copy must exist (a graph whose values were only ever on the host)
    if (!g->device_values_valid)
      if (gs_status s = upload_values(g)) return s;
    // one pass: validate and convert into the pinned stage; nothing of the graph is touched before every value passed
    if (gs_status s = ensure_stage(g, std::max<size_t>(n, 1))) return s;
    double4* tmp = g->h_stage;
    std::atomic<size_t> bad(n);
    gs::parallel_for(n, [&bad, tmp, xyt](size_t b, size_t e) {
      for (size_t l = b; l < e; ++l) {
        const double x = xyt[3 * l], y = xyt[3 * l + 1], th = xyt[3 * l + 2];
        if (!(std::isfinite(x) && std::isfinite(y) && std::isfinite(th))) {
          size_t cur = bad.load();
          while (l < cur && !bad.compare_exchange_weak(cur, l)) {
          }
          return;
        }
        tmp[l] = make_double4(x, y, std::cos(th), std::sin(th));
      }
    });
    // several ranks: the call is collective, so a rank with a bad value still takes part in the exchange below with
    // its old rows and reports afterwards
    const bool ok = bad.load() >= n;
    if (!ok && g->world == 1) return g->fail(GS_EINVAL, "non-finite pose in owned row %zu", bad.load());
    if (ok) {
      if (n) GS_CUDA(g, cudaMemcpyAsync(g->d_pose.p, tmp, sizeof(double4) * n, cudaMemcpyHostToDevice, g->stream));
      g->h2d_bytes += sizeof(double4) * (double)n;
    }
    // ghost rows: the neighbours' new values, device to device
    if (gs_status s = halo_exchange(g, reinterpret_cast<double*>(g->d_pose.p), 4)) return s;
    GS_CUDA(g, cudaStreamSynchronize(g->stream));
    g->host_values_fresh = false;   // the host copy is stale on every rank (gs_get_poses / gs_prepare bring it back)
    g->device_values_valid = true;
    g->system_valid = false;
    if (!ok) return g->fail(GS_EINVAL, "non-finite pose in owned row %zu (the rank's rows were left as they were)", bad.load());
  }
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_set_poses(gs_graph* g, size_t n, const uint64_t* keys, const double* xyt) try {
  if (!g || (n && (!keys || !xyt))) return GS_EINVAL;
  if (gs_status s = download_values(g)) return s;
  // validate everything before touching the values: an unknown key or a non-finite pose leaves the graph as it was
  {
    std::atomic<size_t> bad_key(n), bad_val(n);   // lowest offending index of each kind
    gs::parallel_for(n, [&](size_t b, size_t e) {
      for (size_t k = b; k < e; ++k) {
        if (g->find(keys[k]) < 0) {
          size_t cur = bad_key.load();
          while (k < cur && !bad_key.compare_exchange_weak(cur, k)) {
          }
          return;
        }
        if (!(std::isfinite(xyt[3 * k]) && std::isfinite(xyt[3 * k + 1]) && std::isfinite(xyt[3 * k + 2]))) {
          size_t cur = bad_val.load();
          while (k < cur && !bad_val.compare_exchange_weak(cur, k)) {
          }
          return;
        }
      }
    });
    if (bad_key.load() < n) return g->fail(GS_ENOKEY, "unknown key %llu", (unsigned long long)keys[bad_key.load()]);
    if (bad_val.load() < n) return g->fail(GS_EINVAL, "non-finite pose for key %llu", (unsigned long long)keys[bad_val.load()]);
  }
  gs::parallel_for(n, [g, keys, xyt](size_t b, size_t e) {
    for (size_t k = b; k < e; ++k) {
      const int v = g->find(keys[k]);
      g->xyt[3 * v] = xyt[3 * k];
      g->xyt[3 * v + 1] = xyt[3 * k + 1];
      g->xyt[3 * v + 2] = xyt[3 * k + 2];
    }
  });
  g->host_values_fresh = true;
  g->theta_wrapped = false;
  g->device_values_valid = false;
  g->system_valid = false;
  if (!g->topo_dirty) return upload_values(g);
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_chi2(gs_graph* g, double* chi2) try {
  if (!g || !chi2) return GS_EINVAL;
  if (gs_status s = prepare(g, nullptr)) return s;
  if (gs_status s = launch_chi2(g, g->d_pose.p)) return s;
  if (gs_status s = read_scalars(g)) return s;
  *chi2 = g->h_sc->chi2;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_num_vertices(const gs_graph* g, size_t* n) try {
  if (!g || !n) return GS_EINVAL;
  *n = g->keys.size();
  return GS_OK;
} catch (...) {
  return on_exception(const_cast<gs_graph*>(g));
}

gs_status gs_num_factors(const gs_graph* g, size_t* n) try {
  if (!g || !n) return GS_EINVAL;
  *n = g->e1.size() + g->pr_v.size();
  return GS_OK;
} catch (...) {
  return on_exception(const_cast<gs_graph*>(g));
}

gs_status gs_get_transfer_bytes(const gs_graph* g, double* h2d, double* d2h) try {
  if (!g) return GS_EINVAL;
  if (h2d) *h2d = g->h2d_bytes;
  if (d2h) *d2h = g->d2h_bytes;
  return GS_OK;
} catch (...) {
  return on_exception(const_cast<gs_graph*>(g));
}

gs_status gs_get_pattern(gs_graph* g, int32_t* rowptr, int32_t* colidx, size_t* nnzb) try {
  if (!g) return GS_EINVAL;
  if (gs_status s = prepare(g, nullptr)) return s;
  std::vector<int32_t> rp, ci;
  gs::global_pattern(g->N, g->E, g->di.data(), g->dj.data(), rp, ci);
  if (nnzb) *nnzb = ci.size();
  if (rowptr) std::memcpy(rowptr, rp.data(), sizeof(int32_t) * rp.size());
  if (colidx) std::memcpy(colidx, ci.data(), sizeof(int32_t) * ci.size());
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_get_system(gs_graph* g, double* blocks, double* b, double* chi2) try {
  if (!g) return GS_EINVAL;
  if (g->world != 1) return g->fail(GS_EINVAL, "gs_get_system is a single-rank parity hook");
  if (gs_status s = prepare(g, nullptr)) return s;
  if (gs_status s = launch_assemble(g)) return s;
  if (gs_status s = read_scalars(g)) return s;
  if (chi2) *chi2 = g->h_sc->chi2;
  const int N = g->N;
  const size_t hs = g->hstride;
  if (b) {
    GS_CUDA(g, cudaMemcpyAsync(b, g->d_b.p, sizeof(double) * 3 * N, cudaMemcpyDeviceToHost, g->stream));
    GS_CUDA(g, cudaStreamSynchronize(g->stream));
  }
  if (blocks) {
    std::vector<double> ho(9 * hs), hd(6 * (size_t)N);
    GS_CUDA(g, cudaMemcpyAsync(ho.data(), g->d_Ho.p, sizeof(double) * 9 * hs, cudaMemcpyDeviceToHost, g->stream));
    GS_CUDA(g, cudaMemcpyAsync(hd.data(), g->d_Hd.p, sizeof(double) * 6 * N, cudaMemcpyDeviceToHost, g->stream));
    GS_CUDA(g, cudaStreamSynchronize(g->stream));
    size_t w = 0;
    const LocalStructure& S = g->S;
    auto put_off = [&](int word) {
      const int pos = word >> 1;
      const bool tr = word & 1;
      for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) blocks[9 * w + 3 * r + c] = gs::h_entry(ho.data(), hs, pos, r, c, tr);
      ++w;
    };
    for (int r = 0; r < N; ++r) {
      const int sl = r / gs::kSlice, wd = S.rsell.width[sl];
      int s = 0;
      auto col_at = [&](int k) { return S.r_col[S.rsell.pos(r, k)]; };
      while (s < wd && col_at(s) >= 0 && col_at(s) < r) put_off(S.r_blk[S.rsell.pos(r, s++)]);
      const double d00 = hd[r], d01 = hd[(size_t)N + r], d02 = hd[2 * (size_t)N + r];
      const double d11 = hd[3 * (size_t)N + r], d12 = hd[4 * (size_t)N + r], d22 = hd[5 * (size_t)N + r];
      const double full[9] = {d00, d01, d02, d01, d11, d12, d02, d12, d22};
      std::memcpy(blocks + 9 * w, full, sizeof full);
      ++w;
      while (s < wd && col_at(s) >= 0) put_off(S.r_blk[S.rsell.pos(r, s++)]);
    }
  }
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_solve_linear(gs_graph* g, double lambda, double* delta, int* pcg_iters, double* rel_residual) try {
  if (!g) return GS_EINVAL;
  if (g->world != 1 && delta) return g->fail(GS_EINVAL, "gs_solve_linear returns delta on a single rank only");
  if (gs_status s = prepare(g, nullptr)) return s;
  if (!g->system_valid)
    if (gs_status s = launch_assemble(g)) return s;
  int done = 0;
  if (gs_status s = pcg_solve(g, lambda, pcg_iters, rel_residual, &done)) return s;
  if (delta) {
    GS_CUDA(g, cudaMemcpyAsync(delta, g->d_x.p, sizeof(double) * 3 * g->N, cudaMemcpyDeviceToHost, g->stream));
    GS_CUDA(g, cudaStreamSynchronize(g->stream));
  }
  if (done == 2) return g->fail(GS_ENAN, "PCG breakdown (p^T A p <= 0 or NaN)");
  if (done == 3) return g->fail(GS_ENOTCONVERGED, "PCG reached pcg_max_iters");
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

// Marginals(graph, values).marginalCovariance(key): the (key, key) 3x3 block of (J^T Omega J)^-1 at the
// current values, in the tangent space of that pose (the reference has it commented out,
// graph.cpp:120,126-127).  Three PCG solves with unit right-hand sides per pose, undamped, on the
// assembled system; on several ranks the owner of the pose broadcasts the three rows.
gs_status gs_marginal_covariances(gs_graph* g, size_t n, const uint64_t* keys, double* cov9s) try {
  if (!g || (n && (!keys || !cov9s))) return GS_EINVAL;
  for (size_t k = 0; k < n; ++k)
    if (g->find(keys[k]) < 0) return g->fail(GS_ENOKEY, "unknown key %llu", (unsigned long long)keys[k]);
  if (gs_status s = prepare(g, nullptr)) return s;
  if (!g->system_valid)
    if (gs_status s = launch_assemble(g)) return s;
  const size_t nb = sizeof(double) * 3 * (size_t)std::max(g->n_own, 1);
  DevBuf<double> keep, row;
  GS_CUDA(g, keep.alloc(3 * (size_t)std::max(g->n_own, 1)));
  GS_CUDA(g, row.alloc(4));
  GS_CUDA(g, cudaMemcpyAsync(keep.p, g->d_b.p, nb, cudaMemcpyDeviceToDevice, g->stream));
  g->system_valid = false;   // b is borrowed for the unit right-hand sides; valid again once it is back
  gs_status rc = GS_OK;
  const double minus_one = -1.0;
  for (size_t k = 0; k < n && rc == GS_OK; ++k) {
    const int dense = g->ins2dense[g->find(keys[k])];
    const int owner = g->world > 1 ? g->S.part.owner(dense) : 0;
    const int local = dense - (g->world > 1 ? g->S.part.row_begin(g->rank) : 0);
    double cols[3][3];
    for (int c = 0; c < 3 && rc == GS_OK; ++c) {
      GS_CUDA(g, cudaMemsetAsync(g->d_b.p, 0, nb, g->stream));
      if (owner == g->rank)
        GS_CUDA(g, cudaMemcpyAsync(g->d_b.p + 3 * (size_t)local + c, &minus_one, sizeof(double), cudaMemcpyHostToDevice,
                                   g->stream));
      int done = 0;
      rc = pcg_solve(g, 0.0, nullptr, nullptr, &done);
      if (rc != GS_OK) break;
      if (done == 2) rc = g->fail(GS_ENAN, "PCG breakdown while solving for a marginal (is the graph anchored by a prior?)");
      if (done == 3) rc = g->fail(GS_ENOTCONVERGED, "PCG reached pcg_max_iters while solving for a marginal");
      if (rc != GS_OK) break;
      const double* src = g->d_x.p + 3 * (size_t)(owner == g->rank ? local : 0);
      if (g->world > 1) {
        if (ncclResult_t r = gs::nccl().Broadcast(src, row.p, 3, ncclDouble, owner, g->comm, g->stream); r != ncclSuccess) {
          rc = g->fail(GS_ENCCL, "ncclBroadcast: %s", gs::nccl().GetErrorString(r));
          break;
        }
        src = row.p;
      }
      if (cudaMemcpyAsync(cols[c], src, 3 * sizeof(double), cudaMemcpyDeviceToHost, g->stream) != cudaSuccess ||
          cudaStreamSynchronize(g->stream) != cudaSuccess) {
        rc = g->fail(GS_ECUDA, "reading a marginal column: %s", cudaGetErrorString(cudaGetLastError()));
        break;
      }
      g->d2h_bytes += 24.0;
    }
    if (rc != GS_OK) break;
    double* out = cov9s + 9 * k;
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) out[3 * a + b] = 0.5 * (cols[b][a] + cols[a][b]);   // PCG columns agree to its tolerance
  }
  // put the gradient back whatever happened
  if (cudaMemcpyAsync(g->d_b.p, keep.p, nb, cudaMemcpyDeviceToDevice, g->stream) == cudaSuccess &&
      cudaStreamSynchronize(g->stream) == cudaSuccess)
    g->system_valid = true;
  return rc;
} catch (...) {
  return on_exception(g);
}

gs_status gs_marginal_covariance(gs_graph* g, uint64_t key, double cov9[9]) try {
  return gs_marginal_covariances(g, 1, &key, cov9);
} catch (...) {
  return on_exception(g);
}

gs_status gs_get_trace(gs_graph* g, size_t cap, double* errors, double* lambdas, int32_t* tries, size_t* n) try {
  if (!g) return GS_EINVAL;
  const size_t m = g->tr_tries.size();
  if (n) *n = m;
  if (!errors && !lambdas && !tries) return GS_OK;
  if (cap < m) return g->fail(GS_EINVAL, "capacity %zu < %zu outer iterations", cap, m);
  if (errors) std::copy(g->tr_err.begin(), g->tr_err.end(), errors);
  if (lambdas) {
    std::fill(lambdas, lambdas + m, 0.0);
    std::copy(g->tr_lambda.begin(), g->tr_lambda.end(), lambdas);
  }
  if (tries) std::copy(g->tr_tries.begin(), g->tr_tries.end(), tries);
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_profile_pcg(gs_graph* g, double lambda, int iters, double ms[3], double bytes[3]) try {
  if (!g || !ms || iters <= 0) return GS_EINVAL;
  if (gs_status s = prepare(g, nullptr)) return s;
  if (!g->system_valid)
    if (gs_status s = launch_assemble(g)) return s;
  Scalars* h = g->h_sc;
  std::memset(h, 0, sizeof(Scalars));
  h->lambda = lambda;
  h->tol2 = 0.0;  // never converge: every profiled iteration does full work
  h->max_iters = 1 << 30;
  if (coarse_active(g))
    if (gs_status s = coarse_setup(g, lambda)) return s;
  if (gs_status s = write_scalars(g)) return s;
  if (gs_status s = launch_pcg_setup(g)) return s;
  if (gs_status s = begin_p2p_solve(g)) return s;
  for (int w = 0; w < 3; ++w)
    if (gs_status s = launch_pcg_iteration(g, (unsigned int)w)) return s;  // warm-up
  std::vector<cudaEvent_t> ev(4 * (size_t)iters);
  for (auto& e : ev) GS_CUDA(g, cudaEventCreate(&e));
  for (int it = 0; it < iters; ++it) {
    cudaEvent_t* e = &ev[4 * (size_t)it];
    for (int stage = 0; stage < 3; ++stage) {
      cudaEventRecord(e[stage], g->stream);
      if (gs_status s = launch_pcg_stage(g, stage, (unsigned int)(3 + it))) return s;
    }
    cudaEventRecord(e[3], g->stream);
  }
  if (g->p2p) gs::k_p2p_advance<<<1, 32, 0, g->stream>>>(g->d_seq.p, (unsigned int)(3 + iters));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  GS_CUDA(g, cudaGetLastError());
  double acc[3] = {0, 0, 0};
  for (int it = 0; it < iters; ++it)
    for (int k = 0; k < 3; ++k) {
      float t = 0.f;
      cudaEventElapsedTime(&t, ev[4 * (size_t)it + k], ev[4 * (size_t)it + k + 1]);
      acc[k] += t;
    }
  for (auto& e : ev) cudaEventDestroy(e);
  for (int k = 0; k < 3; ++k) ms[k] = acc[k] / iters;
  if (bytes) {
    const double N = g->n_own, nnzb = (double)g->S.nnzo_own + g->n_own;
    bytes[0] = 76.0 * nnzb + 52.0 * N;                 // K3
    bytes[1] = (4 * 24.0 + 72.0 + 3 * 24.0) * N;       // K4a: x,p,r,Ap + Minv read; x,r,z written
    bytes[2] = 72.0 * N;                               // K4b: z,p read; p written
  }
  g->launches += launches_per_pcg_iteration(g) * (int64_t)(iters + 3) + 2;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

gs_status gs_profile_outer(gs_graph* g, int reps, double ms[3], double bytes[3]) try {
  if (!g || !ms || reps <= 0) return GS_EINVAL;
  if (gs_status s = prepare(g, nullptr)) return s;
  GS_CUDA(g, cudaMemsetAsync(g->d_x.p, 0, sizeof(double) * 3 * std::max(g->n_loc, 1), g->stream));
  // The whole sequence (one warm-up pass, then `reps` passes of assemble -> retract -> chi2) is enqueued without a
  // host synchronisation, with an event between the stages — as gs_profile_pcg does — so that a launch's duration
  // is what the stream saw, not the host's launch latency (which is a third of a 13-us k_retract), and each kernel
  // still starts on a cache its predecessors have overwritten (the three working sets together exceed L2).
  std::vector<cudaEvent_t> ev(4 * (size_t)reps);
  for (auto& e : ev) GS_CUDA(g, cudaEventCreate(&e));
  gs_status rc = GS_OK;
  for (int r = -1; r < reps && rc == GS_OK; ++r) {  // r = -1 is the warm-up pass
    cudaEvent_t* e = r >= 0 ? &ev[4 * (size_t)r] : nullptr;
    if (e) cudaEventRecord(e[0], g->stream);
    rc = launch_assemble(g);
    if (rc != GS_OK) break;
    if (e) cudaEventRecord(e[1], g->stream);
    gs::k_retract<<<g->grid_rows, gs::kThreads, 0, g->stream>>>(g->d_pose.p, g->d_x.p, g->d_pose_try.p, g->n_own);
    rc = halo_exchange(g, reinterpret_cast<double*>(g->d_pose_try.p), 4);
    if (rc != GS_OK) break;
    if (e) cudaEventRecord(e[2], g->stream);
    rc = launch_chi2(g, g->d_pose_try.p);
    if (rc != GS_OK) break;
    if (e) cudaEventRecord(e[3], g->stream);
  }
  cudaError_t ce = cudaStreamSynchronize(g->stream);
  if (ce == cudaSuccess) ce = cudaGetLastError();
  double acc[3] = {0, 0, 0};
  if (rc == GS_OK && ce == cudaSuccess)
    for (int r = 0; r < reps; ++r)
      for (int k = 0; k < 3; ++k) {
        float t = 0.f;
        cudaEventElapsedTime(&t, ev[4 * (size_t)r + k], ev[4 * (size_t)r + k + 1]);
        acc[k] += t;
      }
  for (auto& e : ev) cudaEventDestroy(e);
  if (rc != GS_OK) return rc;
  GS_CUDA(g, ce);
  for (int k = 0; k < 3; ++k) ms[k] = acc[k] / reps;
  if (bytes) {
    const double N = g->n_own, E = g->e_count, nnzb = (double)g->S.nnzo_own + g->n_own;
    bytes[0] = 80.0 * E + 24.0 * N + 72.0 * nnzb + 24.0 * N;   // K1+K2
    bytes[1] = 72.0 * N;                                        // K5 retract
    bytes[2] = 80.0 * E;                                        // K5 chi2
  }
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

const char* gs_last_error(const gs_graph* g) { return g ? g->err.c_str() : "null graph"; }

void gs_compose_reference(const double a[3], const double b[3], double out[3]) {
  // graph.hpp:30-43, restated with both quirks
  const double cos_th = std::cos(a[2]);
  const double sin_th = std::sin(a[2]);
  const double dx = b[0], dy = b[1], dth = b[2];
  const double x = (cos_th * dx) + (-sin_th * dy);
  const double y = (sin_th * dx) + (cos_th * dy);
  double th = a[2] + dth;
  th = std::fmod(th + M_PI, 2 * M_PI) - M_PI;
  out[0] = x;
  out[1] = y;
  out[2] = th;
}

void gs_compose(const double a[3], const double b[3], double out[3]) {
  const double c = std::cos(a[2]), s = std::sin(a[2]);
  const double x = a[0] + c * b[0] - s * b[1];
  const double y = a[1] + s * b[0] + c * b[1];
  const double th = a[2] + b[2];
  out[0] = x;
  out[1] = y;
  out[2] = std::atan2(std::sin(th), std::cos(th));
}

gs_status gs_closest_keyframe(gs_graph* g, uint64_t key, int skip, uint64_t* out_key, double* out_dist) try {
  if (!g || !out_key || skip < 0) return GS_EINVAL;
  const int v = g->find(key);
  if (v < 0) return g->fail(GS_ENOKEY, "unknown key %llu", (unsigned long long)key);
  const int nv = (int)g->keys.size();
  if (nv <= skip) return g->fail(GS_ENOKEY, "not enough keyframes (%d <= skip %d)", nv, skip);
  if (gs_status s = prepare(g, nullptr)) return s;
  const int count = nv - skip;
  // candidates in insertion order -> dense rows (identity when the keys were inserted in increasing order)
  const int* order = nullptr;
  if (!g->identity_order) {
    if (g->closest_order_n != nv) {
      GS_CUDA(g, g->d_closest_order.alloc(nv));
      GS_CUDA(g, cudaMemcpyAsync(g->d_closest_order.p, g->ins2dense.data(), sizeof(int) * nv, cudaMemcpyHostToDevice, g->stream));
      g->h2d_bytes += sizeof(int) * (double)nv;
      g->closest_order_n = nv;
    }
    order = g->d_closest_order.p;
  }
  const int grid = (count + gs::kThreads * gs::kClosestPerThread - 1) / (gs::kThreads * gs::kClosestPerThread);
  GS_CUDA(g, g->d_closest.alloc((size_t)grid + 2 + gs::kMaxRanks));
  gs::ClosestOut* part = g->d_closest.p + 2 + gs::kMaxRanks;
  gs::ClosestOut* res = g->d_closest.p;                      // [0] this rank's result, [2 ...] all ranks' results
  // the query pose stays on the device: its owner's row, broadcast to the other ranks
  const int qdense = g->ins2dense[v];
  const double4* query = nullptr;
  if (g->world == 1) {
    query = g->d_pose.p + qdense;
  } else {
    const int owner = g->S.part.owner(qdense);
    double* qbuf = reinterpret_cast<double*>(g->d_pose_all.p);   // scratch: only the final pose gather uses it
    const double* src = reinterpret_cast<const double*>(g->d_pose.p + (owner == g->rank ? qdense - g->S.v0 : 0));
    GS_NCCL(g, gs::nccl().Broadcast(src, qbuf, 4, ncclDouble, owner, g->comm, g->stream));
    query = reinterpret_cast<const double4*>(qbuf);
  }
  gs::k_closest<<<grid, gs::kThreads, 0, g->stream>>>(g->d_pose.p, order, count, query, g->world > 1 ? g->S.v0 : 0, g->n_own, part,
                                                       g->d_counter.p + 5, res);
  g->launches += 1;
  gs::ClosestOut all[gs::kMaxRanks];
  if (g->world > 1) {
    GS_NCCL(g, gs::nccl().AllGather(res, res + 2, sizeof(gs::ClosestOut), ncclChar, g->comm, g->stream));
    GS_CUDA(g, cudaMemcpyAsync(all, res + 2, sizeof(gs::ClosestOut) * g->world, cudaMemcpyDeviceToHost, g->stream));
  } else {
    GS_CUDA(g, cudaMemcpyAsync(all, res, sizeof(gs::ClosestOut), cudaMemcpyDeviceToHost, g->stream));
  }
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  g->d2h_bytes += sizeof(gs::ClosestOut) * (double)g->world;
  double bd = all[0].dist;
  int bi = all[0].idx;
  for (int q = 1; q < g->world; ++q)
    if (all[q].dist < bd || (all[q].dist == bd && all[q].idx < bi)) {
      bd = all[q].dist;
      bi = all[q].idx;
    }
  if (bi < 0 || bi >= nv) return g->fail(GS_ECUDA, "closest_keyframe: no candidate found");
  *out_key = g->keys[bi];
  if (out_dist) *out_dist = bd;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

// ---- closest_keyframe without any allocation ---------------------------------------------------------------------
// gs_closest_keyframe keeps its per-CTA partial results, its result word and its last-CTA counter in a device buffer
// that it allocates (and grows) on demand.  Callers that must not allocate — code under CUDA-graph capture, tight
// serving loops, harnesses that forbid cudaMalloc — get the same kernel with that scratch in a file-scope device
// array of fixed size instead.  Every GPU that loads the library holds the array (64 KB + 32 B) until the process
// exits.  The array's name is a host-side stand-in in host code, not a device address: the address is resolved once
// per device with cudaGetSymbolAddress, with that device current, and handed to the kernel.
constexpr int kClosestStaticMaxCtas = 4096;                                                 // partials the array holds
constexpr int kClosestStaticMaxPoses = kClosestStaticMaxCtas * gs::kThreads * gs::kClosestPerThread;   // 8,388,608
static_assert(kClosestStaticMaxPoses == GS_CLOSEST_NOALLOC_MAX_POSES, "header and scratch array disagree");
// [0] result, [1] last-CTA counter (first word), [2 ...] one partial per CTA
static __device__ __align__(256) gs::ClosestOut g_closest_static[2 + kClosestStaticMaxCtas];

namespace {
struct ClosestStaticAddr {
  std::mutex mu;
  gs::ClosestOut* ptr[64] = {nullptr};
};
ClosestStaticAddr& closest_static_addr() {
  static ClosestStaticAddr a;
  return a;
}
}  // namespace

gs_status gs_closest_keyframe_noalloc(gs_graph* g, uint64_t key, int skip, uint64_t* out_key, double* out_dist) try {
  if (!g || !out_key || skip < 0) return GS_EINVAL;
  if (g->world != 1) return g->fail(GS_EINVAL, "gs_closest_keyframe_noalloc runs on a single-rank graph");
  const int v = g->find(key);
  if (v < 0) return g->fail(GS_ENOKEY, "unknown key %llu", (unsigned long long)key);
  const int nv = (int)g->keys.size();
  if (nv <= skip) return g->fail(GS_ENOKEY, "not enough keyframes (%d <= skip %d)", nv, skip);
  // everything that would allocate must have happened before: structure built, values on the device
  if (g->topo_dirty || !g->device_values_valid)
    return g->fail(GS_EINVAL, "gs_closest_keyframe_noalloc needs a prepared graph with its values on the device (call gs_prepare "
                              "or gs_optimize after the last gs_add_* / gs_set_poses): it allocates nothing itself");
  const int count = nv - skip;
  if (count > kClosestStaticMaxPoses)
    return g->fail(GS_EINVAL, "gs_closest_keyframe_noalloc handles at most %d candidate keyframes (%d asked for): its scratch is a "
                              "fixed device array; use gs_closest_keyframe", kClosestStaticMaxPoses, count);
  const int* order = nullptr;
  if (!g->identity_order) {   // the insertion-order -> row table is uploaded (and allocated) by gs_closest_keyframe only
    if (g->closest_order_n != nv)
      return g->fail(GS_EINVAL, "gs_closest_keyframe_noalloc: keys were not inserted in increasing order and the order table is not "
                                "on the device yet; call gs_closest_keyframe once after the last topology change");
    order = g->d_closest_order.p;
  }
  int dev = 0;
  GS_CUDA(g, cudaGetDevice(&dev));
  if (g->prm.device >= 0 && dev != g->prm.device) {
    GS_CUDA(g, cudaSetDevice(g->prm.device));
    dev = g->prm.device;
  }
  if (dev < 0 || dev >= 64) return g->fail(GS_EINVAL, "device ordinal %d out of range", dev);
  gs::ClosestOut* scratch = nullptr;
  {
    ClosestStaticAddr& A = closest_static_addr();
    std::lock_guard<std::mutex> lock(A.mu);
    if (!A.ptr[dev]) {   // once per device, with that device current
      void* p = nullptr;
      GS_CUDA(g, cudaGetSymbolAddress(&p, g_closest_static));
      A.ptr[dev] = static_cast<gs::ClosestOut*>(p);
    }
    scratch = A.ptr[dev];
  }
  const int grid = (count + gs::kThreads * gs::kClosestPerThread - 1) / (gs::kThreads * gs::kClosestPerThread);
  // the array keeps the last call's contents: the counter must start at zero (a call that failed mid-way could have
  // left it otherwise); result and partials are fully overwritten before they are read
  GS_CUDA(g, cudaMemsetAsync(scratch + 1, 0, sizeof(gs::ClosestOut), g->stream));
  gs::k_closest<<<grid, gs::kThreads, 0, g->stream>>>(g->d_pose.p, order, count, g->d_pose.p + g->ins2dense[v], 0, g->n_own,
                                                       scratch + 2, reinterpret_cast<unsigned int*>(scratch + 1), scratch);
  g->launches += 1;
  gs::ClosestOut res;
  GS_CUDA(g, cudaMemcpyAsync(&res, scratch, sizeof res, cudaMemcpyDeviceToHost, g->stream));
  GS_CUDA(g, cudaStreamSynchronize(g->stream));
  GS_CUDA(g, cudaGetLastError());
  g->d2h_bytes += sizeof res;
  if (res.idx < 0 || res.idx >= nv) return g->fail(GS_ECUDA, "closest_keyframe: no candidate found");
  *out_key = g->keys[res.idx];
  if (out_dist) *out_dist = res.dist;
  return GS_OK;
} catch (...) {
  return on_exception(g);
}

}  // extern "C"
