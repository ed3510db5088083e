// Multi-GPU PCG iteration with the collectives fused into the kernels (NVLink peer memory, no NCCL
// call and no extra launch inside the iteration).
//
// Every rank exports one small "window" allocation (cudaIpc) that holds its flags, the landing zone
// for the segment sums of all ranks, and the ghost rows of its direction vector p.  Peers map it and
// write into it directly (the owned part of p stays in ordinary memory: keeping the whole vector in
// the exported allocation made every kernel that streams p several times slower):
//
//   k_spmv_p2p    waits until its neighbours have pushed the ghost rows of p for this iteration,
//                 multiplies, and its last CTA stores this rank's p.Ap segment sums into EVERY
//                 rank's window, fences, and raises a flag there.
//   k_update_p2p  waits for all ranks' p.Ap flags, adds the 8 segment sums in the fixed global order
//                 (so alpha has the same bits everywhere), updates x, r, z and pushes its r.z / r.r
//                 segment sums the same way.
//   k_dir_p2p     waits for all r.z / r.r flags, forms beta and the stopping test, updates p and
//                 stores the rows its neighbours need straight into their ghost slots; the last CTA
//                 raises the "p ready" flag at the neighbours.
//
// Flags carry the global iteration sequence number, which is a kernel argument (position inside the
// captured batch) plus a device-resident base that only moves between batches, so no kernel reads a
// word that another CTA of the same launch writes.  Every spin is bounded: after ~1 s a rank gives
// up, marks the solve failed (done = 4) and every later kernel returns at once.
#pragma once
#include "kernels.cuh"

namespace gs {

constexpr int kMaxRanks = kSegments;
constexpr size_t kWindowHeaderBytes = 4096;   // flags + segment sums; the ghost rows of p start here

struct P2PWindow {
  unsigned int halo_flag[kMaxRanks];   // [source rank] p pushed by that rank is valid for spmv number (value - 1)
  unsigned int pap_flag[kMaxRanks];    // [source rank] its p.Ap segment sums of iteration (value - 1) have landed
  unsigned int rz_flag[kMaxRanks];
  unsigned int pad[kMaxRanks];
  double seg_pap[2][kSegments];        // [iteration parity][global segment]
  double seg_rz[2][kSegments];
  double seg_rr[2][kSegments];
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
};

struct HaloPush {
  const int* tile_ptr;         // per tile: range of entries
  const int* row;              // local row to send
  double* const* dst;          // address of its 3 doubles in the peer's window
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

// What one CTA learns at the top of a kernel: whether the wait succeeded, and the totals it waited for.
struct WaitOut {
  int ok;
  double tot[2];
};

// Lane q of warp 0 spins until flags[q] >= target (q in mask); false on timeout (then sc->done = 4).
// The same warp then reads the NT landed segment-sum arrays ONCE per CTA (lanes 0..7, one request)
// and adds them in the fixed global order; everybody else takes the totals from shared memory.
// (Letting every thread read the eight sums past L1 put ~10^5 requests per launch on a single L2
// slice and made the three kernels 3-10x slower than their single-GPU forms.)
template <int NT>
__device__ __forceinline__ bool wait_flags(const unsigned int* flags, unsigned int mask, unsigned int target, Scalars* sc,
                                           long long limit, WaitOut* out, int which, const double* seg0,
                                           const double* seg1) {
  if (threadIdx.x < 32) {
    int ok = 1;
    const long long w0 = clock64();
    if (threadIdx.x < kMaxRanks && ((mask >> threadIdx.x) & 1u)) {
      const long long t0 = clock64();
      while ((int)(ld_relaxed(flags + threadIdx.x) - target) < 0) {
        if (clock64() - t0 > limit) {
          ok = 0;
          break;
        }
      }
    }
    ok = __all_sync(0xffffffffu, ok);
    __threadfence();
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      const double* seg = t == 0 ? seg0 : seg1;
      const double v = threadIdx.x < kSegments ? __ldcg(seg + threadIdx.x) : 0.0;
      double acc = 0.0;
#pragma unroll
      for (int g = 0; g < kSegments; ++g) acc += __shfl_sync(0xffffffffu, v, g);
      if (threadIdx.x == 0) out->tot[t] = acc;
    }
    if (threadIdx.x == 0) {
      if (!ok) sc->done = 4;
      out->ok = ok;
      if (blockIdx.x == 0) {
        sc->spin[which] += clock64() - w0;
        if (which == 0) sc->spin[3] += 1;
      }
    }
  }
  __syncthreads();
  return out->ok != 0;
}

// tile partials -> this rank's segment sums -> every rank's window, then the flag.  NQ quantities,
// dst_of(window, q) gives the landing array of quantity q.
template <int NQ, typename FlagOf, typename DstOf>
__device__ __forceinline__ void reduce_push(const ReduceCtx& c, const P2PCtx& x, const double (&v)[NQ], unsigned int par,
                                            unsigned int stamp, FlagOf flag_of, DstOf dst_of, double* red, int* flag) {
  __shared__ double mysegs[NQ > 1 ? NQ : 1][kSegments];
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const double s = block_sum(v[q], red);
    if (threadIdx.x == 0) c.part[(size_t)q * c.part_stride + blockIdx.x] = s;
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
    // thread q serves rank q: data, system fence, flag
    if ((int)threadIdx.x < x.world) {
      P2PWindow* wq = x.win[threadIdx.x];
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        double* dst = dst_of(wq, q) + (size_t)par * kSegments + (size_t)x.rank * x.spr;
        for (int s = 0; s < x.spr; ++s) dst[s] = mysegs[q][s];
      }
      __threadfence_system();
      st_release_sys(flag_of(wq) + x.rank, stamp);
    }
    if (threadIdx.x == 0) *c.counter = 0u;
  }
}

__global__ void __launch_bounds__(kThreads) k_spmv_p2p(const SpmvView a, const double* __restrict__ p, const double* __restrict__ pg,
                                                        double* __restrict__ y, const ReduceCtx c, const P2PCtx x) {
  __shared__ double red[8];
  __shared__ int flag;
  __shared__ WaitOut wo;
  if (c.sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  if (!wait_flags<0>(me->halo_flag, x.nbr_mask, S + 1, c.sc, x.spin_limit, &wo, 0, nullptr, nullptr)) return;
  const double lambda = c.sc->lambda;
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[1] = {0.0};
  if (r < a.n_own) v[0] = spmv_row(a, p, pg, y, r, lambda);
  reduce_push<1>(c, x, v, S & 1u, S + 1, [](P2PWindow* w) { return w->pap_flag; },
                 [](P2PWindow* w, int) { return &w->seg_pap[0][0]; }, red, &flag);
}

__global__ void __launch_bounds__(kThreads) k_update_p2p(const PcgView a, const ReduceCtx c, const P2PCtx x) {
  __shared__ double red[8];
  __shared__ int flag;
  __shared__ WaitOut wo;
  if (c.sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  if (!wait_flags<1>(me->pap_flag, (1u << x.world) - 1u, S + 1, c.sc, x.spin_limit, &wo, 1, me->seg_pap[S & 1u], nullptr))
    return;
  const double pAp = wo.tot[0];
  if (!(pAp > 0.0)) {  // breakdown: the same on every rank, nobody goes on
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      c.sc->pAp = pAp;
      c.sc->done = 2;
    }
    return;
  }
  const double alpha = c.sc->rzv[(S + 1u) & 1u] / pAp;   // r.z of the previous iterate
  const int r = blockIdx.x * kThreads + threadIdx.x;
  double v[2] = {0.0, 0.0};
  if (r < a.n_own) pcg_update_row(a, r, alpha, v[0], v[1]);
  reduce_push<2>(c, x, v, S & 1u, S + 1, [](P2PWindow* w) { return w->rz_flag; },
                 [](P2PWindow* w, int q) { return q == 0 ? &w->seg_rz[0][0] : &w->seg_rr[0][0]; }, red, &flag);
}

// p = z + beta p ; rows that neighbours need go straight into their ghost slots
__global__ void __launch_bounds__(kThreads) k_dir_p2p(const PcgView a, const ReduceCtx c, const P2PCtx x, const HaloPush h) {
  __shared__ int flag;
  __shared__ WaitOut wo;
  Scalars* sc = c.sc;
  if (sc->done) return;
  const unsigned int S = *x.seq_base + x.k_local;
  P2PWindow* me = x.me;
  if (!wait_flags<2>(me->rz_flag, (1u << x.world) - 1u, S + 1, sc, x.spin_limit, &wo, 2, me->seg_rz[S & 1u],
                     me->seg_rr[S & 1u]))
    return;
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
  const int r = blockIdx.x * kThreads + threadIdx.x;
  if (r < a.n_own) pcg_dir_row(a, r, beta);
  __syncthreads();
  const int e0 = h.tile_ptr[blockIdx.x], e1 = h.tile_ptr[blockIdx.x + 1];
  bool pushed = false;
  for (int e = e0 + (int)threadIdx.x; e < e1; e += kThreads) {
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
    flag = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (flag) {
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
