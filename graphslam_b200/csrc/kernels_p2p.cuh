// Multi-GPU PCG iteration with the collectives fused into the kernels (NVLink peer memory, no NCCL
// call and no extra launch inside the iteration).
//
// Every rank exports one small "window" allocation (cudaIpc) that holds its flags, the landing zone
// for the segment sums of all ranks, and the ghost rows of its direction vector p.  Peers map it and
// write into it directly (the owned part of p stays in ordinary memory: keeping the whole vector in
// the exported allocation made every kernel that streams p several times slower):
//
//   k_spmv_p2p    tiles that multiply ghost rows wait until the neighbours have pushed them for this
//                 iteration (interior tiles start at once); the last CTA sends this rank's p.Ap segment
//                 sums to EVERY rank's window as flagged words.
//   k_update_p2p  one warp per CTA spins on the 16 flagged words of p.Ap, rebuilds the 8 segment sums and
//                 adds them in the fixed global order (so alpha has the same bits everywhere); the CTA
//                 updates x, r, z and the last CTA sends r.z / r.r the same way.
//   k_dir_p2p     waits for the 32 words of r.z / r.r, forms beta and the stopping test, updates p and
//                 stores the rows its neighbours need straight into their ghost slots; the last SENDING
//                 tile raises the "p ready" flag at the neighbours (sending tiles are scheduled first).
//
// Stamps and flags carry the global iteration sequence number, which is a kernel argument (position inside
// the captured batch) plus a device-resident base that only moves between batches, so no kernel reads a
// word that another CTA of the same launch writes.  Every spin is bounded: after ~1 s a rank gives up,
// marks the solve failed (done = 4) and every later kernel returns at once.
//
// Measured on 2 x B200, C3 (500k rows per rank), per iteration: 206 us with every thread reading the landed
// sums -> 111 us (one warp per CTA reads them) -> 99 us (flagged words instead of data + system fence +
// flag) -> 93 us (sending tiles first, only they take the ticket); the three kernels now cost what the
// single-GPU kernels cost on a 500k-row graph, except the direction update (18 us against 10 us).
#pragma once
#include "kernels.cuh"
#include "kernels_coarse.cuh"

namespace gs {

constexpr int kMaxRanks = kSegments;
constexpr size_t kWindowHeaderBytes = 4096;   // flags + segment sums; the ghost rows of p start here

// Segment sums travel as "flagged words" (the low-latency protocol NCCL calls LL): every double is split into
// two 8-byte words {32 data bits, 32-bit iteration stamp}.  An aligned 8-byte store is delivered whole, so a
// reader that sees the stamp has the data: no fence on the writer, no separate flag, one round trip on the reader.
constexpr int kLLWordsPerQuantity = 2 * kSegments;   // 16

struct P2PWindow {
  unsigned int halo_flag[kMaxRanks];   // [source rank] p pushed by that rank is valid for spmv number (value - 1)
  unsigned int pad[24];                // keeps the word arrays on their own 128-byte lines
  unsigned long long ll_pap[2][kLLWordsPerQuantity];       // [iteration parity][2 * global segment + half]
  unsigned long long ll_upd[2][2 * kLLWordsPerQuantity];   // r.z words, then r.r words
  unsigned long long ll_cmu[2][kLLWordsPerQuantity];       // two-level preconditioner: c.mu segment sums
};
static_assert(sizeof(P2PWindow) <= kWindowHeaderBytes, "window header too small");

struct P2PCtx {
  P2PWindow* me;               // the local window
  P2PWindow* const* win;       // device array [rank]: the local window and the peer mappings (a by-value array here
                               // would be indexed dynamically and push the whole argument block onto the stack)
  int world, rank, spr;
  unsigned int nbr_mask;       // ranks this rank exchanges halo rows with
  const unsigned int* seq_base;  // device word: iterations completed before the current batch
  unsigned int k_local;        // position of this iteration inside the batch
  unsigned int* ticket;        // last-CTA counter for the halo push
  long long spin_limit;        // clock64 ticks before a wait gives up
  const int* tile_halo;        // per tile: 0, or the 1-based ordinal among the tiles whose rows multiply ghost rows of p
};

struct HaloPush {
  const int* row;              // local row to send
  double* const* dst;          // address of its 3 doubles in the peer's window
  int n_push_tiles;            // tiles with at least one entry: the last of them to finish raises the flags
  const int4* order;           // CTA -> (tile, first entry, end entry, -), sending tiles first: their remote stores and
                               // system fences overlap the rest of the kernel instead of forming its tail (they are
                               // the first and last tiles of the range)
};

// Flags and landing zones live in THIS GPU's memory and are written by peers over NVLink; remote stores
// become visible at this GPU's L2.  The poll is a relaxed L2 read; one gpu-scope acquire fence after
// success orders the data reads behind it (the data itself is read past L1: __ldcg, or lines this
// kernel has not touched before).  A system-scope acquire per poll was measured ~8x slower.
__device__ __forceinline__ unsigned int ld_relaxed(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long ld_relaxed64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// What one CTA learns at the top of a kernel: whether the wait succeeded, and the totals it waited for.
struct WaitOut {
  int ok;
  double tot[2];
};

// Halo flags: lane q of warp 0 spins until flags[q] >= target (q in mask); false on timeout (sc->done = 4).
__device__ __forceinline__ bool wait_flags(const unsigned int* flags, unsigned int mask, unsigned int target, Scalars* sc,
                                           long long limit, WaitOut* out) {
  if (threadIdx.x < 32) {
    int ok = 1;
    const long long w0 = clock64();
    if (threadIdx.x < kMaxRanks && ((mask >> threadIdx.x) & 1u)) {
      while ((int)(ld_relaxed(flags + threadIdx.x) - target) < 0) {
        if (clock64() - w0 > limit) {
          ok = 0;
          break;
        }
      }
    }
    ok = __all_sync(0xffffffffu, ok);
    __threadfence();
    if (threadIdx.x == 0) {
      if (!ok) sc->done = 4;
      out->ok = ok;
    }
  }
  __syncthreads();
  return out->ok != 0;
}

// Segment sums: lane l of warp 0 spins on flagged word l until it carries `stamp`; the warp then rebuilds the
// 8 doubles of each of the NQ quantities and adds them in the fixed global segment order.  Once per CTA: one
// request for the 128/256 bytes (every thread reading them made ~10^5 requests per launch to one L2 slice).
template <int NQ>
__device__ __forceinline__ bool wait_sums(const unsigned long long* words, unsigned int stamp, Scalars* sc, long long limit,
                                          WaitOut* out, int which) {
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    int ok = 1;
    unsigned long long w = 0;
    const long long w0 = clock64();
    if (lane < NQ * kLLWordsPerQuantity) {
      for (;;) {
        w = ld_relaxed64(words + lane);
        if ((unsigned int)(w >> 32) == stamp) break;
        if (clock64() - w0 > limit) {
          ok = 0;
          break;
        }
      }
    }
    ok = __all_sync(0xffffffffu, ok);
    const int half = (int)(unsigned int)(w & 0xffffffffull);
#pragma unroll
    for (int t = 0; t < NQ; ++t) {
      double acc = 0.0;
#pragma unroll
      for (int g = 0; g < kSegments; ++g) {
        const int lo = __shfl_sync(0xffffffffu, half, t * kLLWordsPerQuantity + 2 * g);
        const int hi = __shfl_sync(0xffffffffu, half, t * kLLWordsPerQuantity + 2 * g + 1);
        acc += __hiloint2double(hi, lo);
      }
      if (lane == 0) out->tot[t] = acc;
    }
    if (lane == 0) {
      if (!ok) sc->done = 4;
      out->ok = ok;
      if (blockIdx.x == 0) {
        sc->spin[which] += clock64() - w0;
        if (which == 1) sc->spin[3] += 1;
      }
    }
  }
  __syncthreads();
  return out->ok != 0;
}

// tile partials -> this rank's segment sums -> flagged words in every rank's window (no fence, no flag).
// words_of(window) = the landing array of this reduction; quantity q occupies words [16 q, 16 q + 16) of a parity.
template <int NQ, typename WordsOf>
__device__ __forceinline__ void reduce_push(const ReduceCtx& c, const P2PCtx& x, const double (&v)[NQ], unsigned int par,
                                            unsigned int stamp, WordsOf words_of, double* red, int* flag,
                                            int tile = (int)blockIdx.x) {
  __shared__ double mysegs[NQ > 1 ? NQ : 1][kSegments];
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const double s = block_sum(v[q], red);
    if (threadIdx.x == 0) c.part[(size_t)q * c.part_stride + tile] = s;
  }
  if (last_block(c.counter, flag)) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    for (int s = w; s < c.spr; s += (int)(blockDim.x >> 5)) {
      const int t0 = c.seg_tile[s], t1 = c.seg_tile[s + 1];
      double acc[NQ];
#pragma unroll
      for (int q = 0; q < NQ; ++q) acc[q] = 0.0;
      for (int i = t0 + l; i < t1; i += 32) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) acc[q] += __ldcg(c.part + (size_t)q * c.part_stride + i);
      }
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        const double t = warp_sum(acc[q]);
        if (l == 0) mysegs[q][s] = t;
      }
    }
    __syncthreads();
    // one thread per (destination rank, quantity, segment, half)
    const int per_rank = NQ * x.spr * 2;
    for (int t = threadIdx.x; t < x.world * per_rank; t += blockDim.x) {
      const int d = t / per_rank, u = t - d * per_rank;
      const int q = u / (x.spr * 2), sh = u - q * (x.spr * 2);
      const int s = sh >> 1, h = sh & 1;
      const double val = mysegs[q][s];
      const unsigned int bits = (unsigned int)(h ? __double2hiint(val) : __double2loint(val));
      unsigned long long* dst =
          words_of(x.win[d]) + (size_t)par * (NQ * kLLWordsPerQuantity) + q * kLLWordsPerQuantity + 2 * (x.rank * x.spr + s) + h;
      st_relaxed64(dst, ((unsigned long long)stamp << 32) | bits);
    }
    if (threadIdx.x == 0) *c.counter = 0u;
  }
}

__global__ void __launch_bounds__(kThreads, GS_SPMV_MINB) k_spmv_p2p(const SpmvView a, const double* __restrict__ p, const double* __restrict__ pg,
                                                        double* __restrict__ y, const ReduceCtx c, const P2PCtx x) {
  __shared__ double red[8];
  __shared__ int flag;
  __shared__ WaitOut wo;
  if (c.sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  // only tiles that touch ghost rows wait for the neighbours' pushes: the interior tiles start at once and
  // hide the arrival of the halo (the p.Ap flag below is raised after ALL tiles, so the iteration order holds)
  if (x.tile_halo[blockIdx.x] && !wait_flags(me->halo_flag, x.nbr_mask, S + 1, c.sc, x.spin_limit, &wo)) return;
  const double lambda = c.sc->lambda;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < a.n_own) v[0] = spmv_row(a, p, pg, y, r, lambda);
  reduce_push<1>(c, x, v, S & 1u, S + 1, [](P2PWindow* w) { return &w->ll_pap[0][0]; }, red, &flag);
}

__global__ void __launch_bounds__(kThreads) k_update_p2p(const PcgView a, const ReduceCtx c, const P2PCtx x) {
  __shared__ double red[8];
  __shared__ int flag;
  __shared__ WaitOut wo;
  // (loading the row's 18 operands before the wait was measured slower: 36-40 us against 31-37 us)
  const int r = blockIdx.x * kThreads + threadIdx.x;
  if (c.sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  if (!wait_sums<1>(me->ll_pap[S & 1u], S + 1, c.sc, x.spin_limit, &wo, 1)) return;
  const double pAp = wo.tot[0];
  if (!(pAp > 0.0)) {  // breakdown: the same on every rank, nobody goes on
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      c.sc->pAp = pAp;
      c.sc->done = 2;
    }
    return;
  }
  const double alpha = c.sc->rzv[(S + 1u) & 1u] / pAp;   // r.z of the previous iterate
  double v[2] = {0.0, 0.0};
  if (r < a.n_own) pcg_update_row(a, r, alpha, v[0], v[1]);
  reduce_push<2>(c, x, v, S & 1u, S + 1, [](P2PWindow* w) { return &w->ll_upd[0][0]; }, red, &flag);
}

// p = z + beta p over the tile's 768 contiguous doubles; rows that neighbours need go straight into their ghost slots
__global__ void __launch_bounds__(kThreads) k_dir_p2p(const PcgView a, const ReduceCtx c, const P2PCtx x, const HaloPush h) {
  __shared__ int flag;
  __shared__ WaitOut wo;
  Scalars* sc = c.sc;
  // operands first (they do not depend on beta), then the wait
  const size_t n3 = 3 * (size_t)a.n_own;
  const int4 job = ldg(h.order + blockIdx.x);
  const size_t e0 = (size_t)job.x * (3 * kThreads) + threadIdx.x;
  double zv[3], pv[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const size_t e = e0 + (size_t)k * kThreads;
    zv[k] = e < n3 ? a.z[e] : 0.0;
    pv[k] = e < n3 ? a.p[e] : 0.0;
  }
  if (sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  if (!wait_sums<2>(me->ll_upd[S & 1u], S + 1, sc, x.spin_limit, &wo, 2)) return;
  const double rz = wo.tot[0];
  const double rr = wo.tot[1];
  const double rz_old = sc->rzv[(S + 1u) & 1u];
  const double beta = rz / rz_old;
  const int it = (int)(S - sc->seq0) + 1;
  int done = 0;
  if (rr <= sc->tol2 * sc->rr0) done = 1;
  else if (!(rr == rr)) done = 2;
  else if (it >= sc->max_iters) done = 3;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    sc->rzv[S & 1u] = rz;
    sc->rz = rz;
    sc->rr = rr;
    sc->beta = beta;
    sc->iters = it;
    if (done) sc->done = done;
  }
  if (done) return;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const size_t e = e0 + (size_t)k * kThreads;
    if (e < n3) a.p[e] = dir_value(zv[k], beta, pv[k]);
  }
  const int h0 = job.y, h1 = job.z;
  if (h1 == h0) return;   // nothing to send from this tile (uniform over the CTA)
  __syncthreads();        // the tile's new p is visible to the CTA
  bool pushed = false;
  for (int e = h0 + (int)threadIdx.x; e < h1; e += kThreads) {
    const double* src = a.p + 3 * (size_t)h.row[e];
    double* dst = h.dst[e];
    dst[0] = __ldcg(src);
    dst[1] = __ldcg(src + 1);
    dst[2] = __ldcg(src + 2);
    pushed = true;
  }
  if (pushed) __threadfence_system();   // only the few threads that stored to a peer pay for the system fence
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(x.ticket, 1u);
    flag = (t == (unsigned int)h.n_push_tiles - 1u);
  }
  __syncthreads();
  if (flag) {   // the last sending tile: every neighbour's rows are in place
    if ((int)threadIdx.x < x.world && ((x.nbr_mask >> threadIdx.x) & 1u))
      st_release_sys(&x.win[threadIdx.x]->halo_flag[x.rank], S + 2);
    if (threadIdx.x == 0) *x.ticket = 0u;
  }
}

// ---- two kernels per iteration -------------------------------------------------------------------------------
// The direction update does not need a kernel (nor a cross-GPU signal) of its own: the product can form
// p(c) = z(c) + beta p_prev(c) on the fly for every column it touches, as the single-GPU persistent kernel
// does.  What neighbours then need from each other is z, which the update kernel has just computed:
//
//   k_spmvdir_p2p   waits for r.z / r.r of the previous iteration (beta, stopping test), tiles that touch ghost
//                   columns also for the neighbours' z rows; stores its rows of the new p into the other p
//                   buffer (one designated tile does the same for the ghost rows), multiplies, sends p.Ap.
//   k_updpush_p2p   waits for p.Ap; updates x, r, z; sending tiles (scheduled first) store the z rows their
//                   neighbours need into the neighbours' windows; sends r.z / r.r.
//
// Same arithmetic per row and per reduction as the three-kernel form, hence the same bits.
struct DirSpmvArgs {
  const double* z;        // owned rows
  const double* zg;       // ghost rows of z: in this rank's window, written by the neighbours
  const double* p_prev;   // previous direction, owned rows ...
  const double* pg_prev;  // ... and ghost rows
  double* p_next;
  double* pg_next;
  double* y;              // Ap
  int n_ghost;
  int ghost_chunk;        // ghost doubles advanced by each halo tile: tile_halo[tile] = 1-based ordinal k takes
                          // [(k-1) chunk, k chunk) — spread over the tiles that wait for the halo anyway (one tile
                          // doing all of them put ~10 us on the 8-GPU critical path)
};

__global__ void __launch_bounds__(kThreads, GS_SPMVDIR_MINB) k_spmvdir_p2p(const SpmvView a, const DirSpmvArgs d, const ReduceCtx c,
                                                           const P2PCtx x) {
  __shared__ double red[8];
  __shared__ int flag;
  __shared__ WaitOut wo;
  Scalars* sc = c.sc;
  if (sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  double beta = 0.0;
  if (S != sc->seq0) {   // not the first iteration of the solve: close iteration S-1 (what k_dir_p2p did)
    if (!wait_sums<2>(me->ll_upd[(S - 1u) & 1u], S, sc, x.spin_limit, &wo, 2)) return;
    const double rz = wo.tot[0], rr = wo.tot[1];
    const double rz_old = sc->rzv[S & 1u];
    beta = rz / rz_old;
    const int it = (int)(S - sc->seq0);
    int done = 0;
    if (rr <= sc->tol2 * sc->rr0) done = 1;
    else if (!(rr == rr)) done = 2;
    else if (it >= sc->max_iters) done = 3;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      sc->rzv[(S - 1u) & 1u] = rz;
      sc->rz = rz;
      sc->rr = rr;
      sc->beta = beta;
      sc->iters = it;
      if (done) sc->done = done;
    }
    if (done) return;
  }
  const int halo_ord = x.tile_halo[blockIdx.x];
  if (halo_ord && !wait_flags(me->halo_flag, x.nbr_mask, S + 1, sc, x.spin_limit, &wo)) return;
  if (halo_ord) {   // this tile's share of the ghost rows of the new direction (the next iteration's p_prev)
    const int e0 = (halo_ord - 1) * d.ghost_chunk, e1 = min(e0 + d.ghost_chunk, 3 * d.n_ghost);
    for (int e = e0 + (int)threadIdx.x; e < e1; e += kThreads) d.pg_next[e] = dir_value(d.zg[e], beta, d.pg_prev[e]);
  }
  const double lambda = sc->lambda;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < a.n_own) v[0] = spmv_row_t<true>(a, d.p_prev, d.pg_prev, d.z, d.zg, beta, d.p_next, d.y, r, lambda);
  reduce_push<1>(c, x, v, S & 1u, S + 1, [](P2PWindow* w) { return &w->ll_pap[0][0]; }, red, &flag);
}

__global__ void __launch_bounds__(kThreads) k_updpush_p2p(const PcgView a, const ReduceCtx c, const P2PCtx x, const HaloPush h) {
  __shared__ double red[8];
  __shared__ int flag;
  __shared__ WaitOut wo;
  if (c.sc->done) return;
  const int4 job = ldg(h.order + blockIdx.x);   // (tile, first entry, end entry): sending tiles first
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  if (!wait_sums<1>(me->ll_pap[S & 1u], S + 1, c.sc, x.spin_limit, &wo, 1)) return;
  const double pAp = wo.tot[0];
  if (!(pAp > 0.0)) {  // breakdown: the same on every rank, nobody goes on
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      c.sc->pAp = pAp;
      c.sc->done = 2;
    }
    return;
  }
  const double alpha = c.sc->rzv[(S + 1u) & 1u] / pAp;   // r.z of the previous iterate
  const int r = job.x * kThreads + threadIdx.x;
  double v[2] = {0.0, 0.0};
  if (r < a.n_own) pcg_update_row(a, r, alpha, v[0], v[1]);
  if (job.z > job.y) {   // a sending tile (uniform over the CTA): its new z rows go to the neighbours' ghost slots
    __syncthreads();
    bool pushed = false;
    for (int e = job.y + (int)threadIdx.x; e < job.z; e += kThreads) {
      const double* src = a.z + 3 * (size_t)h.row[e];
      double* dst = h.dst[e];
      dst[0] = __ldcg(src);
      dst[1] = __ldcg(src + 1);
      dst[2] = __ldcg(src + 2);
      pushed = true;
    }
    if (pushed) __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned int t = atomicAdd(x.ticket, 1u);
      flag = (t == (unsigned int)h.n_push_tiles - 1u);
    }
    __syncthreads();
    if (flag) {
      if ((int)threadIdx.x < x.world && ((x.nbr_mask >> threadIdx.x) & 1u))
        st_release_sys(&x.win[threadIdx.x]->halo_flag[x.rank], S + 2);
      if (threadIdx.x == 0) *x.ticket = 0u;
    }
    __syncthreads();
  }
  reduce_push<2>(c, x, v, S & 1u, S + 1, [](P2PWindow* w) { return &w->ll_upd[0][0]; }, red, &flag, job.x);
}

// ---- two-level preconditioner over peer memory ----------------------------------------------------------------
// The iteration of kernels_coarse.cuh with its exchanges fused into the kernels (six launches, no NCCL call):
//
//   k_spmv_p2p          as above (waits for the neighbours' rows of p, sends p.Ap)
//   k_update_p2p        as above (waits for p.Ap; x, r, z = M^-1 r; sends r.M^-1 r and r.r)
//   k_restrict_p2p      c_a = W_a^T r for the rank's own aggregates, stored into EVERY rank's window as flagged words
//   k_c_unpack_p2p      waits for all 3 n_agg entries of c (and, in block 0, for r.M^-1 r / r.r: the scalar step
//                       OP_UPDATE_2L) and writes c as plain doubles
//   k_coarse_mv_p2p     mu = E^-1 c for the rows of the rank's OWN aggregates (1/P of the dense product), c.mu summed
//                       per fixed segment and sent to every rank like the other dot products
//   k_dir_coarse_p2p    waits for the c.mu sums, forms r.z and beta, p = (z + W mu) + beta p; rows the neighbours need
//                       go straight into their ghost slots
//
// The landing zone of c sits in the window between the header and the ghost rows: 2 parities x 6 n_agg words, the
// same size on every rank (n_agg is global).
__device__ __forceinline__ unsigned long long* coarse_words(P2PWindow* w) {
  return reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(w) + kWindowHeaderBytes);
}

__global__ void __launch_bounds__(kThreads, 4) k_restrict_p2p(const CoarseView a, const double* __restrict__ r, const Scalars* sc,
                                                               const P2PCtx x) {
  __shared__ double red[8];
  __shared__ double tot[3];
  if (sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  const int g = blockIdx.x;                   // local aggregate
  double s0, s1, s2;
  restrict_sums(a, r, g, s0, s1, s2);
  const double t0 = block_sum(s0, red);
  const double t1 = block_sum(s1, red);
  const double t2 = block_sum(s2, red);
  if (threadIdx.x == 0) {
    tot[0] = t0;
    tot[1] = t1;
    tot[2] = t2;
  }
  __syncthreads();
  // one thread per (destination rank, component, half): an aligned 8-byte store arrives whole
  const int t = threadIdx.x;
  if (t < 6 * x.world) {
    const int d = t / 6, u = t - 6 * d;
    const double val = tot[u >> 1];
    const unsigned int bits = (unsigned int)((u & 1) ? __double2hiint(val) : __double2loint(val));
    unsigned long long* dst = coarse_words(x.win[d]) + (size_t)(S & 1u) * (6 * (size_t)a.n_agg) + 6 * (size_t)(a.a0 + g) + u;
    st_relaxed64(dst, ((unsigned long long)(S + 1) << 32) | bits);
  }
}

__global__ void __launch_bounds__(kThreads) k_c_unpack_p2p(const P2PCtx x, int n_agg, double* __restrict__ cvec, Scalars* sc) {
  __shared__ WaitOut wo;
  if (sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  const unsigned int stamp = S + 1;
  const int i = blockIdx.x * kThreads + threadIdx.x;
  if (i < 3 * n_agg) {
    const unsigned long long* w = coarse_words(x.me) + (size_t)(S & 1u) * (6 * (size_t)n_agg) + 2 * (size_t)i;
    unsigned long long lo = 0, hi = 0;
    const long long w0 = clock64();
    bool ok = true;
    for (;;) {
      lo = ld_relaxed64(w);
      hi = ld_relaxed64(w + 1);
      if ((unsigned int)(lo >> 32) == stamp && (unsigned int)(hi >> 32) == stamp) break;
      if (clock64() - w0 > x.spin_limit) {
        ok = false;
        break;
      }
    }
    if (ok) cvec[i] = __hiloint2double((int)(unsigned int)(hi & 0xffffffffull), (int)(unsigned int)(lo & 0xffffffffull));
    else sc->done = 4;
  }
  if (blockIdx.x == 0) {   // the scalar step the single-GPU update kernel's last CTA takes (OP_UPDATE_2L)
    if (!wait_sums<2>(x.me->ll_upd[S & 1u], stamp, sc, x.spin_limit, &wo, 2)) return;
    if (threadIdx.x == 0) apply_scalars(sc, OP_UPDATE_2L, wo.tot[0], wo.tot[1]);
  }
}

__global__ void __launch_bounds__(kThreads, 4) k_coarse_mv_p2p(const double* __restrict__ Einv, const double* __restrict__ cvec,
                                                                int m, int row0, int nrows, double* __restrict__ mu,
                                                                double* __restrict__ rowprod, const int* __restrict__ seg_row,
                                                                unsigned int* counter, const Scalars* sc, const P2PCtx x) {
  __shared__ double red[8];
  __shared__ double csegs[kSegments];
  __shared__ int flag;
  if (sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  const int s0 = x.rank * x.spr, s1 = s0 + x.spr;
  if (!coarse_mv_rows(Einv, cvec, m, row0, nrows, mu, rowprod, seg_row, s0, s1, csegs, counter, red, &flag)) return;
  // one thread per (destination rank, own segment, half)
  const int per_rank = x.spr * 2;
  for (int t = threadIdx.x; t < x.world * per_rank; t += blockDim.x) {
    const int d = t / per_rank, sh = t - d * per_rank;
    const int s = sh >> 1, hh = sh & 1;
    const double val = csegs[s0 + s];
    const unsigned int bits = (unsigned int)(hh ? __double2hiint(val) : __double2loint(val));
    unsigned long long* dst = &x.win[d]->ll_cmu[S & 1u][2 * (s0 + s) + hh];
    st_relaxed64(dst, ((unsigned long long)(S + 1) << 32) | bits);
  }
  if (threadIdx.x == 0) *counter = 0u;
}

__global__ void __launch_bounds__(kThreads) k_dir_coarse_p2p(const CoarseView a, const double* __restrict__ z,
                                                              const double* __restrict__ mu, double* __restrict__ p,
                                                              Scalars* sc, const P2PCtx x, const HaloPush h) {
  __shared__ int flag;
  __shared__ WaitOut wo;
  if (sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  const int4 job = ldg(h.order + blockIdx.x);   // (tile, first entry, end entry): sending tiles first
  const int r = job.x * kThreads + (int)threadIdx.x;
  // close the preconditioner application: r.z = r.M^-1 r + c.mu (the 8 segment sums of c.mu in global order), beta.
  // r.z of the previous iterate is read from the other parity slot than the one written here.
  if (!wait_sums<1>(x.me->ll_cmu[S & 1u], S + 1, sc, x.spin_limit, &wo, 2)) return;
  const double rz_new = sc->rz_bj + wo.tot[0];
  const double beta = rz_new / sc->rzv[(S + 1u) & 1u];
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    sc->rzv[S & 1u] = rz_new;
    sc->rz = rz_new;
    sc->beta = beta;
  }
  if (r < a.n_own) dir_coarse_row(a, z, mu, p, beta, r);
  const int h0 = job.y, h1 = job.z;
  if (h1 == h0) return;   // nothing to send from this tile (uniform over the CTA)
  __syncthreads();        // the tile's new p is visible to the CTA
  bool pushed = false;
  for (int e = h0 + (int)threadIdx.x; e < h1; e += kThreads) {
    const double* src = p + 3 * (size_t)h.row[e];
    double* dst = h.dst[e];
    dst[0] = __ldcg(src);
    dst[1] = __ldcg(src + 1);
    dst[2] = __ldcg(src + 2);
    pushed = true;
  }
  if (pushed) __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(x.ticket, 1u);
    flag = (t == (unsigned int)h.n_push_tiles - 1u);
  }
  __syncthreads();
  if (flag) {   // the last sending tile: every neighbour's rows are in place
    if ((int)threadIdx.x < x.world && ((x.nbr_mask >> threadIdx.x) & 1u))
      st_release_sys(&x.win[threadIdx.x]->halo_flag[x.rank], S + 2);
    if (threadIdx.x == 0) *x.ticket = 0u;
  }
}

// start of a solve: the ghost rows of p(0) arrived by the ordinary halo exchange, so mark them
// valid for the first spmv of this solve, and remember where the solve's numbering starts
__global__ void k_p2p_begin(P2PWindow* me, const unsigned int* seq_base, Scalars* sc) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const unsigned int S0 = *seq_base;
    for (int q = 0; q < kMaxRanks; ++q) me->halo_flag[q] = S0 + 1;
    sc->seq0 = S0;
    __threadfence_system();
  }
}

// end of a batch
__global__ void k_p2p_advance(unsigned int* seq_base, unsigned int n) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *seq_base += n;
}

}  // namespace gs
