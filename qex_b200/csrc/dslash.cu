// Staggered Dslash parity sweep for sm_100a (HBM-bound fp64 stencil, no tensor cores).
//
// Restates WHAT src/physics/stagD.nim:348-395 (stagD2) computes,
//   r[s] = a r[s] + b x[s] + sum_k [ U_k(s) x(s+h_k) - U_k(s-h_k)^+ x(s-h_k) ],  s of parity p,
// with the 3-phase shape of the reference (start shifts / local / boundary,
// stagD.nim:355-393) mapped to: face-pack kernels + NCCL exchange on a second
// stream, an interior kernel that skips off-rank neighbours, and per-direction
// exterior kernels that add the halo terms.
//
// One thread per output site; all loads are 128-bit (one complex double) and
// coalesced along the site index.  Links are streamed with ld.global.cs, vectors
// go through the read-only path so the 8/16 neighbour re-reads hit L1/L2.
#include "stencil.cuh"
#include <utility>
#include <algorithm>
#include <cstdlib>

namespace qb {


// Expands LAUNCH once with the compile-time constants NL (link classes), RC (bit j: class j is
// read as 12 reals, bit 2+j: as 14 reals) and DOT that match the link set and the request.
#define QB_CASE(KEY, NLV, RCV, DOTV, ...) \
  case KEY: { constexpr int NL = NLV; constexpr int RC = RCV; constexpr bool DOT = DOTV; (void)DOT; __VA_ARGS__; } break;
#define QB_CASE2(NLV, RCV, ...) \
  QB_CASE(((NLV) == 2 ? 32 : 0) | ((RCV) << 1), NLV, RCV, false, __VA_ARGS__) \
  QB_CASE(((NLV) == 2 ? 32 : 0) | ((RCV) << 1) | 1, NLV, RCV, true, __VA_ARGS__)
#define QB_DISPATCH(g, dot, ...)                                                           \
  do {                                                                                     \
    const int nl_ = (g).nlinks / 4;                                                        \
    const int rc_ = (nl_ == 2) ? (g).rc() : ((g).rc() & 5);                                \
    switch ((nl_ == 2 ? 32 : 0) | (rc_ << 1) | ((dot) ? 1 : 0)) {                          \
      QB_CASE2(1, 0, __VA_ARGS__) QB_CASE2(1, 1, __VA_ARGS__) QB_CASE2(1, 4, __VA_ARGS__)  \
      QB_CASE2(2, 0, __VA_ARGS__) QB_CASE2(2, 1, __VA_ARGS__) QB_CASE2(2, 4, __VA_ARGS__)  \
      QB_CASE2(2, 2, __VA_ARGS__) QB_CASE2(2, 3, __VA_ARGS__) QB_CASE2(2, 6, __VA_ARGS__)  \
      QB_CASE2(2, 8, __VA_ARGS__) QB_CASE2(2, 9, __VA_ARGS__) QB_CASE2(2, 12, __VA_ARGS__) \
      default: return fail("sweep: no kernel for this link set");                          \
    }                                                                                      \
  } while (0)

// Residency: 6 blocks of 128 threads per SM (<= 85 registers) measured fastest on B200 -- the
// compiler keeps more 16-byte loads in flight per thread than at the default 9 blocks/SM
// (tools/tune_sweep.py: 32^4 naive 0.0971 ms vs 0.1085 ms, fat+Naik 0.1827 ms vs 0.1909 ms).
constexpr int SWEEP_MINB = 6;

template <int NL, int RC, bool DOT>
__global__ void __launch_bounds__(SWEEP_BLOCK, SWEEP_MINB) k_sweep(const SweepArgs A) {
  // Nothing is done before the wait: an L2 prefetch of the first wave's forward links there (they do not depend on the
  // previous kernel) was measured SLOWER than the plain wait (32^4: 32.25 vs 31.59 ms per solve, profiles/r03_pdl_ab.log).
  pdl_wait();      // x, the done flag and the reduction scratch belong to the kernels before this one
  pdl_trigger();
  if (A.skip != nullptr && *A.skip != 0.0) return;
  const int idx = blockIdx.x * SWEEP_BLOCK + threadIdx.x;
  double dot = 0.0;
  if (idx < A.Vh) {
    const Site s = site_from_idx(idx, A.parity, A.X);
    double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
    stencil_local<NL, RC>(A, s, idx, acc);
    dot = site_epilogue<DOT>(A, idx, acc);
    if (DOT && A.comm_mask) {
      // multi-rank: sites with an off-rank hop are finished (and counted) by k_exterior
      constexpr int dep = (NL == 2) ? 3 : 1;
      bool bnd = false;
#pragma unroll
      for (int mu = 0; mu < 4; mu++)
        if ((A.comm_mask >> mu) & 1) bnd = bnd || s.x[mu] < dep || s.x[mu] >= A.X[mu] - dep;
      if (bnd) dot = 0.0;
    }
  }
  if (DOT) grid_reduce<SWEEP_BLOCK>(dot, A.partials, A.ticket, A.result);
}

#ifdef QEXB200_TUNE
// launch-shape experiments (QEXB200_SWEEP_TUNE): same body, other block size / residency
template <int NL, int BLOCK, int MINB>
__global__ void __launch_bounds__(BLOCK, MINB) k_sweep_t(const SweepArgs A) {
  const int idx = blockIdx.x * BLOCK + threadIdx.x;
  if (idx < A.Vh) {
    const Site s = site_from_idx(idx, A.parity, A.X);
    double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
    stencil_local<NL, 0>(A, s, idx, acc);
    site_epilogue<false>(A, idx, acc);
  }
}
#endif

// ---- exterior: one launch for all split directions -----------------------------------------
// Adds the terms whose neighbour lives on another rank (boundarySB2 of stagD.nim:381-393).
// ghost_f holds the + neighbour's first `depth` layers of x; ghost_b[j] holds U_k^+ x of the
// - neighbour's last h_j layers.  The boundary sites are enumerated as DISJOINT regions
// (region i: interior in split directions m_1..m_{i-1}, boundary in m_i, anything elsewhere),
// one thread per site, which handles every split direction of its site in a fixed order:
// no site is touched by two threads, so corners need no atomics and the sum is deterministic.
template <int NL, int RC, bool DOT>
__global__ void __launch_bounds__(SWEEP_BLOCK) k_exterior(const ExtArgs A) {
  if (A.skip != nullptr && *A.skip != 0.0) return;
  if (A.wait) {
    if ((int)threadIdx.x < 2 * A.ndir) {
      const ExtDir& D = A.d[threadIdx.x >> 1];
      wait_flag((threadIdx.x & 1) ? D.flag_b : D.flag_f, A.seq, A.err, A.timeout_ns);
    }
    __syncthreads();
  }
  const int t = blockIdx.x * SWEEP_BLOCK + threadIdx.x;
  double dot = 0.0;
  if (t < A.reg[A.nreg - 1].end) dot = exterior_site<NL, RC, DOT>(A, t);
  if (DOT) grid_reduce<SWEEP_BLOCK>(dot, A.partials, A.ticket, A.result, A.xr, A.interior);
}

// ---- face packing: one launch for all split directions --------------------------------------
// startSB of stagD.nim:355-361: x of my first `depth` layers for the - neighbour, and U_k^+ x of
// my last h_j layers for the + neighbour.  On the peer-memory path the destinations are the
// neighbours' receive buffers (NVLink stores) and the last block to finish raises the
// neighbours' flags; on the NCCL path they are the local send buffer.
// Called by every thread of every packing block after its stores: the last block to arrive
// raises the neighbours' flags (release at system scope, after the data of all blocks).
__device__ __forceinline__ void pack_signal(const PackArgs& A, unsigned int nblocks) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int tk = atomicAdd(A.ticket, 1u);
    if (tk == nblocks - 1) {
      *A.ticket = 0u;
      __threadfence_system();
      for (int i = 0; i < A.ndir; i++) {
        st_release_sys(A.d[i].flag_f, A.seq);
        st_release_sys(A.d[i].flag_b, A.seq);
      }
    }
  }
}

template <int NL, int RC>
__global__ void __launch_bounds__(SWEEP_BLOCK) k_pack(const PackArgs A) {
  const int t = blockIdx.x * SWEEP_BLOCK + threadIdx.x;
  if (t < A.d[A.ndir - 1].end) pack_item<NL, RC>(A, t);
  if (A.ticket) pack_signal(A, gridDim.x);
}

// ---- the whole multi-rank sweep as ONE kernel (peer-memory path) -----------------------------
// Blocks take a role by index, in the order the hardware dispatches them:
//   [0, npack)       pack the faces straight into the neighbours' windows over NVLink and, when the
//                    last of them is done, raise the neighbours' flags;
//   [npack, +ntiles) one TILE of 128 consecutive checkerboard sites each, in the order of a table
//                    built once per geometry: tiles without an off-rank hop first (the transfer is in
//                    flight meanwhile), tiles that touch a face last.  A face tile waits for exactly
//                    the flags its sites need (bits of the table entry), then runs the same stencil
//                    with ghost data substituted for its off-rank hops.
// Every site is visited by exactly one thread and every thread has a site: the grid is the single-
// rank grid plus the pack blocks (the round-1 kernel ran all Vh threads with the boundary ones idle
// and then a third group of blocks for the boundary: +36 % blocks at 1.2.2.2, 0.109 vs 0.070 ms).
// Pack blocks never wait, face tiles wait only on REMOTE pack blocks: no cycle.  The Re<bx, r>
// reduction (and its sum over ranks) closes the same kernel, so a CG iteration on N GPUs launches
// exactly the kernels it launches on one.
template <int NL, int RC, bool DOT>
__global__ void __launch_bounds__(SWEEP_BLOCK, SWEEP_MINB) k_sweep_mr(const MRArgs F) {
  const SweepArgs& A = F.A;
  if (A.skip != nullptr && *A.skip != 0.0) return;   // past convergence: no pack, no flags, no sum -- on every rank alike
  const int bid = blockIdx.x;
  const bool stamp = F.stamps != nullptr && threadIdx.x == 0;
  if (stamp) atomicMin(F.stamps + ST_START, globaltimer_ns());
  double dot = 0.0;
  if (bid < F.npack) {
    // few, persistent pack blocks: a pack item is one link + one vector in flight, so a block per 128 items holds a
    // residency slot for a full load -> NVLink store -> system fence round trip while moving almost nothing (896
    // such blocks = the whole first wave of the fat+Naik sweep on 2 ranks, 22 us of 193).  Here every pack thread
    // strides over the items (stores are fire-and-forget) and the block pays ONE fence at the end.
    const int end = F.P.d[F.P.ndir - 1].end;
#pragma unroll 1
    for (int t = bid * SWEEP_BLOCK + threadIdx.x; t < end; t += F.npack * SWEEP_BLOCK) pack_item<NL, RC>(F.P, t);
    pack_signal(F.P, F.npack);
    if (stamp) atomicMax(F.stamps + ST_PACK_END, globaltimer_ns());
  } else {
    const unsigned int e = F.order[bid - F.npack];
    const int tile = (int)(e & 0xffffffu);
    const unsigned int need = e >> 24;
    if (need) {
      if (threadIdx.x < 8 && ((need >> threadIdx.x) & 1u)) {
        const unsigned int* fl = nullptr;
#pragma unroll
        for (int k = 0; k < 8; k++)
          if ((int)threadIdx.x == k) fl = (k & 1) ? F.gd[k >> 1].flag_b : F.gd[k >> 1].flag_f;
        if (F.stamps) {
          const unsigned long long w0 = globaltimer_ns();
          const bool ready = ld_acquire_sys(fl) == F.seq;
          wait_flag(fl, F.seq, F.err, F.timeout_ns);
          if (!ready) {
            const unsigned long long dt = globaltimer_ns() - w0;
            atomicMax(F.stamps + ST_WAIT_MAX, dt);
            atomicAdd(F.stamps + ST_WAIT_SUM, dt);
            atomicAdd(F.stamps + ST_WAIT_N, 1ull);
          }
        } else {
          wait_flag(fl, F.seq, F.err, F.timeout_ns);
        }
      }
      __syncthreads();
      if (stamp) atomicMin(F.stamps + ST_FACE_START, globaltimer_ns());
    }
    const int idx = tile * SWEEP_BLOCK + threadIdx.x;
    if (idx < A.Vh) {
      const Site s = site_from_idx(idx, A.parity, A.X);
      double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
      if (need) stencil_ghost<NL, RC>(F, s, idx, acc);
      else stencil_local<NL, RC>(A, s, idx, acc);
      dot = site_epilogue<DOT>(A, idx, acc);
    }
    if (stamp) atomicMax(F.stamps + (need ? ST_FACE_END : ST_INT_END), globaltimer_ns());
  }
  if (DOT) grid_reduce<SWEEP_BLOCK>(dot, A.partials, A.ticket, A.result, F.xr);
}

static inline unsigned int* win_flag(char* win, int buf, int mu, int side) {
  return reinterpret_cast<unsigned int*>(win) + ((buf * 4 + mu) * 2 + side);
}
static inline double2* win_recv(char* win, int buf, size_t cap) {
  return reinterpret_cast<double2*>(win + WIN_HDR) + (size_t)buf * cap;
}

// Launched on `st` (the comm stream).  seq is the sweep's exchange number.
static int fill_pack(const Gauge& g, int q, const double2* x, unsigned int seq, PackArgs& P) {
  Context& c = ctx();
  P.L = g.links(); P.x = x;
  for (int i = 0; i < 4; i++) P.X[i] = c.g.X[i];
  P.Vh = c.g.Vh; P.q = q; P.ndir = (int)c.halo.size();
  P.seq = seq; P.ticket = c.p2p ? c.d_ticket + 1 : nullptr;
  const int nl = g.nlinks / 4;
  const int buf = seq & 1;
  int end = 0;
  for (int i = 0; i < P.ndir; i++) {
    const HaloDir& hd = c.halo[i];
    PackDir& D = P.d[i];
    D.mu = hd.mu; D.depth = hd.depth; D.faceh = hd.faceh;
    end += (hd.depth + (nl == 2 ? 4 : 1)) * hd.faceh;
    D.end = end;
    if (c.p2p) {
      double2* rm = win_recv(c.peer_win[hd.rank_minus], buf, c.halo_cap);
      double2* rp = win_recv(c.peer_win[hd.rank_plus], buf, c.halo_cap);
      D.out_f = rm + hd.off_f;
      D.out_b0 = rp + hd.off_b;
      D.out_b1 = rp + hd.off_b + (size_t)3 * hd.faceh;
      D.flag_f = win_flag(c.peer_win[hd.rank_minus], buf, hd.mu, 0);
      D.flag_b = win_flag(c.peer_win[hd.rank_plus], buf, hd.mu, 1);
    } else {
      D.out_f = c.sendbuf + hd.off_f;
      D.out_b0 = c.sendbuf + hd.off_b;
      D.out_b1 = c.sendbuf + hd.off_b + (size_t)3 * hd.faceh;
      D.flag_f = D.flag_b = nullptr;
    }
#ifdef QEXB200_FAULT_INJECT   // never in the product binary: a deliberately wrong halo (the x layers land in the U^+ x slots and
    std::swap(D.out_f, D.out_b0);   // vice versa) to show that bench.py's preflight turns a broken exchange into exit code 3
#endif
  }
  return end;   // work items
}

// NCCL path: launched on `st` (the comm stream).  seq is the sweep's exchange number.
int pack_faces(const Gauge& g, int q, const double2* x, unsigned int seq, cudaStream_t st) {
  PackArgs P;
  const int end = fill_pack(g, q, x, seq, P);
  const int grid = (end + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
  QB_DISPATCH(g, false, k_pack<NL, RC><<<grid, SWEEP_BLOCK, 0, st>>>(P));
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

// Boundary enumeration + ghost pointers of the current exchange; returns the number of sites.
static int fill_ext(const Gauge& g, int parity, double2* r, double scale, const double2* bx,
                    double* dot_out, const double* dot_interior, ExtArgs& E) {
  Context& c = ctx();
  E.bx = bx; E.partials = c.d_partials; E.ticket = c.d_ticket; E.result = dot_out;
  E.interior = dot_interior; E.xr = xrank();
  E.L = g.links(); E.r = r; E.scale = scale;
  for (int i = 0; i < 4; i++) { E.X[i] = c.g.X[i]; E.dep[i] = 0; E.nbv[i] = 0; }
  E.Vh = c.g.Vh; E.parity = parity; E.ndir = (int)c.halo.size();
  E.seq = c.halo_seq; E.wait = c.p2p ? 1 : 0; E.err = c.d_err;
  E.timeout_ns = c.p2p_timeout_ns;
  const int buf = c.halo_seq & 1;
  const double2* rb = c.p2p ? win_recv(c.win, buf, c.halo_cap) : c.recvbuf;
  for (int i = 0; i < E.ndir; i++) {
    const HaloDir& hd = c.halo[i];
    ExtDir& D = E.d[i];
    const int Xm = c.g.X[hd.mu];
    D.mu = hd.mu; D.depth = hd.depth; D.faceh = hd.faceh;
    D.nb = (2 * hd.depth < Xm) ? 2 * hd.depth : Xm;
    D.ghost_f = rb + hd.off_f;
    D.ghost_b0 = rb + hd.off_b;
    D.ghost_b1 = rb + hd.off_b + (size_t)3 * hd.faceh;
    D.flag_f = c.p2p ? win_flag(c.win, buf, hd.mu, 0) : nullptr;
    D.flag_b = c.p2p ? win_flag(c.win, buf, hd.mu, 1) : nullptr;
    E.dep[hd.mu] = hd.depth; E.nbv[hd.mu] = D.nb;
  }
  return plan_regions(E);
}

// NCCL path.  dot_out != nullptr: also completes Re<bx, r> (interior part already in *dot_interior).
static int exterior(const Gauge& g, int parity, double2* r, double scale, const double2* bx,
                    double* dot_out, const double* dot_interior, const double* skip) {
  Context& c = ctx();
  ExtArgs E;
  const int end = fill_ext(g, parity, r, scale, bx, dot_out, dot_interior, E);
  E.skip = skip;
  if (end == 0) return 0;
  const int grid = (end + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
  QB_DISPATCH(g, dot_out != nullptr, k_exterior<NL, RC, DOT><<<grid, SWEEP_BLOCK, 0, c.stream>>>(E));
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

// Block -> tile table of the multi-rank sweep for output parity `parity` and hop depth of `nl` link classes:
// tiles whose sites all stay on the rank first, then the tiles with a face site, each tagged with the flags
// (bit 2 mu: + neighbour's x layers, bit 2 mu + 1: - neighbour's products) its sites read.  Built once per geometry.
void plan_tiles(const Geom& g, int nl, int parity, std::vector<unsigned int>& out) {
  const int dep = (nl == 2) ? 3 : 1;
  const int ntiles = (g.Vh + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
  std::vector<unsigned int> face;
  out.clear();
  out.reserve((size_t)ntiles);
  const int xh = g.X[0] >> 1;
  for (int t = 0; t < ntiles; t++) {
    unsigned int need = 0;
    const int i1 = (t + 1) * SWEEP_BLOCK < g.Vh ? (t + 1) * SWEEP_BLOCK : g.Vh;
    for (int idx = t * SWEEP_BLOCK; idx < i1; idx++) {
      int r = idx, x[4];
      const int x0h = r % xh; r /= xh;
      x[1] = r % g.X[1]; r /= g.X[1];
      x[2] = r % g.X[2];
      x[3] = r / g.X[2];
      x[0] = 2 * x0h + ((parity + x[1] + x[2] + x[3]) & 1);
      for (int mu = 0; mu < 4; mu++) {
        if (!((g.comm_mask >> mu) & 1)) continue;
        if (x[mu] + dep >= g.X[mu]) need |= 1u << (2 * mu);
        if (x[mu] - dep < 0) need |= 1u << (2 * mu + 1);
      }
    }
    if (need) face.push_back((unsigned int)t | (need << 24));
    else out.push_back((unsigned int)t);
  }
  out.insert(out.end(), face.begin(), face.end());
}

static bool overlap_on() {
  const char* ev = getenv("QEXB200_OVERLAP");
  return !(ev && ev[0] == '0');
}

static int tile_order(int nl, int parity, const unsigned int** order, int* ntiles) {
  Context& c = ctx();
  unsigned int*& d = c.d_order[nl - 1][parity];
  const int n = (c.g.Vh + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
  if (!d) {
    if (n >= (1 << 24)) return fail("sweep: %d tiles exceed the 24-bit tile index", n);
    std::vector<unsigned int> h;
    plan_tiles(c.g, nl, parity, h);
    if (!overlap_on()) {
      // measurement switch (QEXB200_OVERLAP=0): face tiles FIRST, so the whole grid queues up behind the halo wait and
      // nothing hides the transfer -- the "overlap off" arm of the reference's 3-phase protocol (stagD.nim:355-393)
      std::stable_partition(h.begin(), h.end(), [](unsigned int e) { return (e >> 24) != 0; });
    }
    QB_CUDA(cudaMalloc(&d, sizeof(unsigned int) * h.size()));
    QB_CUDA(cudaMemcpy(d, h.data(), sizeof(unsigned int) * h.size(), cudaMemcpyHostToDevice));
  }
  *order = d; *ntiles = n;
  return 0;
}

int sweep(const Gauge& g, int parity, double2* r, const double2* x, const double2* bx, double a,
          double b, double sgn, double post, double* dot_out, const double* skip) {
  Context& c = ctx();
  if (!g.loaded) return fail("sweep: gauge not loaded");
  const bool multi = c.g.comm_mask != 0;

  SweepArgs A;
  A.L = g.links(); A.x = x; A.r = r; A.bx = bx ? bx : r;
  A.a = a; A.b = bx ? b : 0.0; A.sgn = sgn; A.post = post;
  for (int i = 0; i < 4; i++) A.X[i] = c.g.X[i];
  A.Vh = c.g.Vh; A.parity = parity; A.comm_mask = c.g.comm_mask;
  A.partials = c.d_partials; A.ticket = c.d_ticket; A.result = dot_out;
  A.skip = skip;
  const int grid = (c.g.Vh + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
  const bool fused_dot = dot_out != nullptr;

  if (multi && c.p2p) {
    // peer-memory path: pack + all tiles (+ dot over ranks) in one launch
    QB_CHECK(halo_begin(g));
    MRArgs F;
    F.A = A;
    const int npk = fill_pack(g, 1 - parity, x, c.halo_seq, F.P);
    const int nl = g.nlinks / 4;
    QB_CHECK(tile_order(nl, parity, &F.order, &F.ntiles));
    const int buf = c.halo_seq & 1;
    const double2* rb = win_recv(c.win, buf, c.halo_cap);
    for (int mu = 0; mu < 4; mu++) F.gd[mu] = GhostDir{nullptr, nullptr, nullptr, nullptr, nullptr, 0};
    for (const HaloDir& hd : c.halo) {
      GhostDir& D = F.gd[hd.mu];
      D.gf = rb + hd.off_f;
      D.gb0 = rb + hd.off_b;
      D.gb1 = rb + hd.off_b + (size_t)3 * hd.faceh;
      D.flag_f = win_flag(c.win, buf, hd.mu, 0);
      D.flag_b = win_flag(c.win, buf, hd.mu, 1);
      D.faceh = hd.faceh;
    }
    F.npack = std::min((npk + SWEEP_BLOCK - 1) / SWEEP_BLOCK, c.pack_blocks);
    F.seq = c.halo_seq; F.err = c.d_err; F.timeout_ns = c.p2p_timeout_ns; F.xr = xrank();
    F.stamps = c.d_stamps ? c.d_stamps + (size_t)ST_N * (c.halo_seq % STAMP_RING) : nullptr;
    const int total = F.npack + F.ntiles;
    if ((size_t)total > c.npart) return fail("sweep: %d blocks exceed the %zu reduction slots", total, c.npart);
    // residency 6 blocks/SM measured best for every variant in round 1 (5, 4, 3 were 1-20 % slower)
    QB_DISPATCH(g, fused_dot, k_sweep_mr<NL, RC, DOT><<<total, SWEEP_BLOCK, 0, c.stream>>>(F));
    count_launch();
    QB_CUDA(cudaGetLastError());
    return 0;
  }

#ifdef QEXB200_TUNE   // launch-shape experiments (tools/tune_sweep.py), not part of the product binary
  const int tune = getenv("QEXB200_SWEEP_TUNE") ? atoi(getenv("QEXB200_SWEEP_TUNE")) : 0;
  if (!multi && !fused_dot && tune > 0) {
#define QB_T(ID, BLOCK, MINB)                                                               \
    if (tune == ID) {                                                                         \
      const int gr = (c.g.Vh + BLOCK - 1) / BLOCK;                                            \
      if (g.nlinks == 8) k_sweep_t<2, BLOCK, MINB><<<gr, BLOCK, 0, c.stream>>>(A);            \
      else k_sweep_t<1, BLOCK, MINB><<<gr, BLOCK, 0, c.stream>>>(A);                          \
    }
    QB_T(1, 128, 8) QB_T(2, 128, 6) QB_T(3, 128, 4) QB_T(4, 256, 4) QB_T(5, 256, 3) QB_T(6, 64, 16)
    QB_T(7, 64, 12) QB_T(8, 96, 8) QB_T(9, 128, 10) QB_T(10, 256, 2) QB_T(11, 64, 8) QB_T(12, 32, 24)
#undef QB_T
    count_launch();
    QB_CUDA(cudaGetLastError());
    return 0;
  }
#endif
  if (multi) QB_CHECK(halo_start(g, parity, x));
  // NCCL path: the interior part of the dot goes to a scratch scalar, k_exterior finishes it
  double* dot_interior = c.d_scal + 48;
  if (multi) A.result = dot_interior;
  if (multi && !overlap_on()) QB_CHECK(halo_wait());   // measurement switch: exchange first, then the interior
  const bool pdl = c.pdl_loop && !multi;
  QB_DISPATCH(g, fused_dot, QB_CUDA(launch_k(k_sweep<NL, RC, DOT>, dim3(grid), dim3(SWEEP_BLOCK), c.stream, pdl, A)));
  count_launch();
  QB_CUDA(cudaGetLastError());
  if (multi) {
    QB_CHECK(halo_wait());
    QB_CHECK(exterior(g, parity, r, post * sgn, bx, dot_out, dot_interior, skip));
  }
  return 0;
}

}  // namespace qb
