// Device link pipeline: fat7 + Lepage + Naik links (makeImpLinks,
// src/gauge/fat7l.nim:78-162), HISQ (src/physics/hisqLinks.nim:32-39), nHYP
// (src/gauge/hypsmear.nim:49-162), projectU (src/maths/matrixFunctions.nim:301-312) and plaq
// (src/gauge/gaugeUtils.nim:213-284) on the resident link fields.
//
// The reference builds every staple from pre-shifted copies of whole fields (Shifter ^*, three
// extra passes per staple) and projects in separate sweeps.  Here ONE kernel shape does all of it:
// a thread owns a site, gathers the six matrices of each staple pair from the checkerboarded
// fields by coordinate arithmetic, sums up to three weighted pairs onto `cbase*base + acc`,
// optionally projects, and writes once.  nHYP is then 12 + 12 + 4 launches (one per output
// field, projection fused), fat7 is 10 launches per (mu, nu) pair reduced to 4.
// HBM-bound streaming of 144-byte matrices; fp64 FMA pipe is the secondary limit.
// On a lattice split over ranks the fields live in a symmetric peer-mapped arena and a site's off-rank
// neighbours are read straight from the owner's memory over NVLink (linkops.cuh m3_ldx<MR>), with a
// one-warp cross-rank barrier between dependent launches: no packing, no exchange buffers, no NCCL.
#include "linkops.cuh"
#include "stencil.cuh"

namespace qb {

#ifndef QB_SB
#define QB_SB 64
#endif
constexpr int SB = QB_SB;

// ---- pool of matrix-field temporaries (kept across calls: HMC smears every force step) -------
// On a multi-rank context with a peer window the fields come from the symmetric arena (links_alloc), in the
// same order on every rank, so a neighbour's temporary sits at the same arena offset.
struct MatPool {
  std::vector<double2*> free_list;          // LIFO: deterministic across ranks
  std::vector<std::pair<double2*, bool>> all;
  size_t elems = 0;
};
static MatPool g_pool;

int smear_free() {
  for (auto& e : g_pool.all) links_free(e.first, 1, e.second);
  g_pool.all.clear(); g_pool.free_list.clear(); g_pool.elems = 0;
  return 0;
}
static int pool_get(double2** out) {
  Context& c = ctx();
  const size_t n = (size_t)18 * c.g.Vh;
  if (g_pool.elems != n) { smear_free(); g_pool.elems = n; }
  if (!g_pool.free_list.empty()) { *out = g_pool.free_list.back(); g_pool.free_list.pop_back(); return 0; }
  double2* p = nullptr;
  bool arena = false;
  QB_CHECK(links_alloc(1, &p, &arena));
  g_pool.all.push_back(std::make_pair(p, arena));
  *out = p;
  return 0;
}
static void pool_put(double2* p) { if (p) g_pool.free_list.push_back(p); }
int pool_acquire(double2** out) { return pool_get(out); }
void pool_release(double2* p) { pool_put(p); }

// Multi-rank: every launch of the pipeline is preceded by a cross-rank barrier (what it reads remotely was
// written by the peers' previous launches; what it overwrites may still be read by their previous launch).
static inline bool multi() { return ctx().g.nranks > 1; }
static int mr_begin(Peers* pr, std::initializer_list<const void*> fields) {
  if (!multi()) { pr->me = 0; return 0; }
  QB_CHECK(peers_fill(pr));
  for (const void* f : fields)
    if (f && !in_arena(f)) return fail("link pipeline on a split lattice: a link set lives outside the peer arena (create gauges after qexb200_p2p_attach)");
  return 0;
}
struct Tmp {   // RAII handle of one pooled field
  double2* p = nullptr;
  ~Tmp() { pool_put(p); }
  int get() { return p ? 0 : pool_get(&p); }
  void drop() { pool_put(p); p = nullptr; }
};

// ---- the staple kernel ----------------------------------------------------------------------
struct StapleTerm {
  const double2* G;   // side link field (direction dG)
  const double2* L;   // middle link field (direction dL of the launch)
  int dG;
  double coef;
};
struct StapleArgs {
  StapleTerm t[3];
  int dL;
  const double2* base;  // out = cbase*base + acc + sum_k coef_k * pair_k   (base / acc may be null)
  double cbase;
  const double2* acc;   // may alias out
  double2* out;         // projected when PROJ
  double2* save;        // NS == 1: the bare pair of term 0 (computeGenStaple's saved staple), may be null
  double2* raw;         // PROJ: the unprojected sum (hypsmear's l1x/l2x/flx, needed by the force), may be null
  int X[4];
  int Vh;
  Peers pr;             // multi-rank: where the neighbours' copies of the fields are
};

// Blocks alternate parity (block b: parity b & 1, sites (b >> 1) * SB ...): the even and the odd sites of
// one neighbourhood run at the same time and share their reads in L2.  With the two parities in separate
// halves of the grid every field was fetched from HBM twice (ncu: 604 MB read for 302 MB of input).
#ifndef QB_STAPLE_MINB
#define QB_STAPLE_MINB 6
#endif
template <int NS, bool PROJ, bool MR>
__global__ void __launch_bounds__(SB, QB_STAPLE_MINB) k_staple(const StapleArgs a) {
  const int par = blockIdx.x & 1, idx = (blockIdx.x >> 1) * SB + threadIdx.x;
  if (idx >= a.Vh) return;
  const LSite s = lsite(idx, par, a.X);
  M3 r = m3_zero();
  if (a.base) m3_axpy(r, a.cbase, m3_ld(a.base, s, a.Vh));
  if (a.acc) m3_axpy(r, 1.0, m3_ld(a.acc, s, a.Vh));
#pragma unroll
  for (int k = 0; k < NS; k++) {
    const M3 st = staple_pair_x<MR>(a.t[k].G, a.t[k].L, a.t[k].dG, a.dL, s, a.X, a.Vh, a.pr);
    if (NS == 1 && a.save) m3_st(a.save, s, a.Vh, st);
    m3_axpy(r, a.t[k].coef, st);
  }
  if (PROJ) {
    if (a.raw) m3_st(a.raw, s, a.Vh, r);
    r = project_u(r);
  }
  m3_st(a.out, s, a.Vh, r);
}

static inline int grid_par(int Vh, int block) { return 2 * ((Vh + block - 1) / block); }

// ---- the same staple sum with the gathers staged through shared memory (single rank) --------------------
// k_staple above is bound by latency: six 144-byte matrices per pair pass through registers, at 168 registers
// a thread has at most three of them in flight and an SM holds 12 warps (ncu: 18 % of the warp slots, DRAM at
// 35-47 %, FP64 pipe at 46-56 %).  Measured (32^4, profiles/r01_staple_pipe_ab.log): HISQ 16.0 -> 15.1 ms,
// nHYP 5.6-5.8 -> 5.4-5.5 ms; bit-identical results (tests/test_gpu_links.py).  Here a thread streams its sites through a private shared-memory ring with
// cp.async (LDGSTS): the ring has one slot for the upper bracket's three matrices and one for the lower's, and
// each slot is refilled for the NEXT (site, pair) as soon as it has been read, so a full pair of matrix products
// (and the projection, when fused) runs between the request and the use of every matrix.  No block barrier:
// a thread only ever reads what it staged itself.  Slot reuse: the ld.shared of a slot are issued before the
// cp.async that refill it (program order, "memory" clobbers), and a refill cannot land before its global read
// returns (>= an L2 round trip, hundreds of cycles after the 16-byte shared loads have been served in order by the
// same LSU); the bit-for-bit A/B test against the register kernel runs this reuse millions of times per launch.  Blocks are persistent and keep the parity interleave
// (the grid is even, block b stays on parity b & 1).  The arithmetic is that of staple_pair_x, in the same order.
#ifndef QB_PB
#define QB_PB 64
#endif
#ifndef QB_CPASYNC
#define QB_CPASYNC "cp.async.cg.shared.global"   // .cg: past L1 -- with the ring taking the carve-out .ca thrashes what little L1 is left (measured 21.8 vs 15.1 ms)
#endif
constexpr int PB = QB_PB;                   // threads per block
constexpr int PIPE_SMEM = 54 * PB * 16;     // 6 matrices * 9 elements * 16 B per thread = 55296 B per block
#ifndef QB_PIPE_MINB
#define QB_PIPE_MINB 4
#endif

__device__ __forceinline__ void cp16(double2* smem, const double2* g) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile(QB_CPASYNC " [%0], [%1], 16;" ::"r"(sa), "l"(g) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_wait1() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void cp_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// request matrix `f` at site s into ring position m (0..5)
__device__ __forceinline__ void stage_m3(double2* my, int m, const double2* f, const LSite& s, int Vh) {
  const double2* p = f + (size_t)s.par * 9 * Vh + s.idx;
#pragma unroll
  for (int e = 0; e < 9; e++) cp16(my + (m * 9 + e) * PB, p + (size_t)e * Vh);
}
__device__ __forceinline__ M3 ring_m3(const double2* my, int m) {
  M3 r;
#pragma unroll
  for (int e = 0; e < 9; e++) r.e[e] = my[(m * 9 + e) * PB];
  return r;
}
// upper bracket of pair t at site s: G(s), L(s+dG), G(s+dL); lower: G(s-dG), L(s-dG), G(s-dG+dL)
__device__ __forceinline__ void stage_up(double2* my, const StapleTerm& t, int dL, const LSite& s, const int* X, int Vh) {
  stage_m3(my, 0, t.G, s, Vh);
  stage_m3(my, 1, t.L, lhop(s, t.dG, +1, X), Vh);
  stage_m3(my, 2, t.G, lhop(s, dL, +1, X), Vh);
}
__device__ __forceinline__ void stage_down(double2* my, const StapleTerm& t, int dL, const LSite& s, const int* X, int Vh) {
  const LSite sb = lhop(s, t.dG, -1, X);
  stage_m3(my, 3, t.G, sb, Vh);
  stage_m3(my, 4, t.L, sb, Vh);
  stage_m3(my, 5, t.G, lhop(sb, dL, +1, X), Vh);
}

template <int NS, bool PROJ>
__global__ void __launch_bounds__(PB, QB_PIPE_MINB) k_staple_pipe(const StapleArgs a, int ntiles) {
  extern __shared__ double2 ring[];
  double2* my = ring + threadIdx.x;
  const int par = blockIdx.x & 1;           // gridDim.x is even: the parity of a block never changes
  int tile = blockIdx.x;
  int idx = (tile >> 1) * PB + threadIdx.x;
  bool live = tile < ntiles && idx < a.Vh;
  LSite s = lsite(live ? idx : 0, par, a.X);
  if (live) stage_up(my, a.t[0], a.dL, s, a.X, a.Vh);
  cp_commit();
  if (live) stage_down(my, a.t[0], a.dL, s, a.X, a.Vh);
  cp_commit();
  while (tile < ntiles) {                   // uniform over the block
    const int ntile = tile + gridDim.x;
    const int nidx = (ntile >> 1) * PB + threadIdx.x;
    const bool nlive = ntile < ntiles && nidx < a.Vh;
    const LSite ns = lsite(nlive ? nidx : 0, par, a.X);
    // cbase*base or acc (the launcher never passes both here): requested now, first used after the first pair
    const double2* isrc = a.base ? a.base : a.acc;
    const double c0 = a.base ? a.cbase : 1.0;
    M3 ini = m3_zero();
    if (live && isrc) ini = m3_ld(isrc, s, a.Vh);
    M3 r = m3_zero();
#pragma unroll
    for (int k = 0; k < NS; k++) {
      const bool last = k + 1 == NS;
      cp_wait1();                           // the upper bracket of (site, k) has landed; only the lower one may be in flight
      M3 st;
      {
        const M3 A = ring_m3(my, 0), B = ring_m3(my, 1), C = ring_m3(my, 2);
        if (!last) { if (live) stage_up(my, a.t[(k + 1) % NS], a.dL, s, a.X, a.Vh); }
        else if (nlive) stage_up(my, a.t[0], a.dL, ns, a.X, a.Vh);
        cp_commit();
        st = m3_mul_na(m3_mul(A, B), C);
      }
      cp_wait1();                           // the lower bracket has landed; the refill just issued may be in flight
      {
        const M3 A = ring_m3(my, 3), B = ring_m3(my, 4), C = ring_m3(my, 5);
        if (!last) { if (live) stage_down(my, a.t[(k + 1) % NS], a.dL, s, a.X, a.Vh); }
        else if (nlive) stage_down(my, a.t[0], a.dL, ns, a.X, a.Vh);
        cp_commit();
        const M3 lo = m3_mul(m3_mul_an(A, B), C);
        m3_axpy(st, 1.0, lo);
      }
      if (live) {
        if (NS == 1 && a.save) m3_st(a.save, s, a.Vh, st);
        if (k == 0 && isrc) m3_axpy(r, c0, ini);
        m3_axpy(r, a.t[k].coef, st);
      }
    }
    if (live) {
      if (PROJ) {
        if (a.raw) m3_st(a.raw, s, a.Vh, r);
        r = project_u(r);
      }
      m3_st(a.out, s, a.Vh, r);
    }
    tile = ntile; idx = nidx; live = nlive; s = ns;
  }
  cp_wait0();
}

// QEXB200_STAPLE_PIPE=1 selects the shared-memory-staged kernel on a single rank (A/B runs; bit-identical results,
// HISQ 16.0 -> 15.1 ms at 32^4).  OPT-IN: the refill of a ring slot right after its ld.shared is ordered by a latency
// argument, not by a cp.async.wait; the register-gather kernel is the default until racecheck has run on it.
static bool staple_pipe_enabled() {
  const char* e = getenv("QEXB200_STAPLE_PIPE");   // read per launch: the A/B test flips it inside one process
  return e && e[0] == '1';
}
// Persistent grid = what is resident at once (the ring needs the full shared-memory carve-out to fit
// QB_PIPE_MINB blocks per SM; a grid larger than one wave would serialise whole blocks).
template <int NS, bool PROJ>
static int launch_pipe_t(const StapleArgs& a, int ntiles, cudaStream_t st) {
  static int resident = 0, resident_dev = -1;   // function attributes are per device: redo them after init on another one
  int dev = 0;
  QB_CUDA(cudaGetDevice(&dev));
  if (!resident || resident_dev != dev) {
    QB_CUDA(cudaFuncSetAttribute(k_staple_pipe<NS, PROJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, PIPE_SMEM));
    QB_CUDA(cudaFuncSetAttribute(k_staple_pipe<NS, PROJ>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    int sms = 0, occ = 0;
    resident_dev = dev;
    QB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    QB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_staple_pipe<NS, PROJ>, PB, PIPE_SMEM));
    if (occ < 1) return fail("k_staple_pipe does not fit on an SM");
    resident = sms * occ;
    if (getenv("QEXB200_PIPE_VERBOSE")) fprintf(stderr, "k_staple_pipe<%d,%d>: %d blocks/SM, %d resident\n", NS, (int)PROJ, occ, resident);
  }
  int grid = resident;
  if (const char* e = getenv("QEXB200_PIPE_GRID")) grid = atoi(e) > 1 ? atoi(e) : 2;   // tests: few blocks, many tiles each
  grid -= grid & 1;
  if (grid > ntiles) grid = ntiles;
  k_staple_pipe<NS, PROJ><<<grid, PB, PIPE_SMEM, st>>>(a, ntiles);
  return 0;
}
static int launch_staple_pipe(const StapleArgs& a, int ns, bool proj) {
  Context& c = ctx();
  const int ntiles = grid_par(c.g.Vh, PB);
  if (proj) {
    if (ns == 1) return launch_pipe_t<1, true>(a, ntiles, c.stream);
    if (ns == 2) return launch_pipe_t<2, true>(a, ntiles, c.stream);
    return launch_pipe_t<3, true>(a, ntiles, c.stream);
  }
  if (ns == 1) return launch_pipe_t<1, false>(a, ntiles, c.stream);
  if (ns == 2) return launch_pipe_t<2, false>(a, ntiles, c.stream);
  return launch_pipe_t<3, false>(a, ntiles, c.stream);
}
template <bool MR>
static void launch_staple_t(const StapleArgs& a, int ns, bool proj, int grid, cudaStream_t st) {
  if (proj) {
    if (ns == 1) k_staple<1, true, MR><<<grid, SB, 0, st>>>(a);
    else if (ns == 2) k_staple<2, true, MR><<<grid, SB, 0, st>>>(a);
    else k_staple<3, true, MR><<<grid, SB, 0, st>>>(a);
  } else {
    if (ns == 1) k_staple<1, false, MR><<<grid, SB, 0, st>>>(a);
    else if (ns == 2) k_staple<2, false, MR><<<grid, SB, 0, st>>>(a);
    else k_staple<3, false, MR><<<grid, SB, 0, st>>>(a);
  }
}
static int launch_staple(const StapleArgs& a, int ns, bool proj) {
  Context& c = ctx();
  const int grid = grid_par(c.g.Vh, SB);
  if (multi()) {
    for (int k = 0; k < ns; k++)
      if (!in_arena(a.t[k].G) || !in_arena(a.t[k].L)) return fail("link pipeline on a split lattice: a field lives outside the peer arena");
    QB_CHECK(xbarrier());
    launch_staple_t<true>(a, ns, proj, grid, c.stream);
  } else if (staple_pipe_enabled() && !(a.base && a.acc)) {
    QB_CHECK(launch_staple_pipe(a, ns, proj));
  } else {
    launch_staple_t<false>(a, ns, proj, grid, c.stream);
  }
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}
static Peers g_peers = {};   // filled by mr_begin at the start of every pipeline entry point
static StapleArgs staple_args(int dL) {
  StapleArgs a = {};
  const Geom& g = ctx().g;
  for (int d = 0; d < 4; d++) a.X[d] = g.X[d];
  a.Vh = g.Vh; a.dL = dL;
  a.pr = g_peers;
  return a;
}

// out = projectU(in) / out = c * in, n matrix fields back to back
template <bool PROJ>
__global__ void __launch_bounds__(SB) k_sitewise(double2* out, const double2* in, double c, int Vh, size_t nsites) {
  const size_t t = (size_t)blockIdx.x * SB + threadIdx.x;
  if (t >= nsites) return;
  const size_t blk = t / Vh;            // (field, parity) block of 9*Vh
  const int idx = (int)(t - blk * Vh);
  const double2* p = in + blk * 9 * Vh + idx;
  M3 m;
#pragma unroll
  for (int e = 0; e < 9; e++) m.e[e] = p[(size_t)e * Vh];
  if (PROJ) m = project_u(m);
  else {
#pragma unroll
    for (int e = 0; e < 9; e++) { m.e[e].x *= c; m.e[e].y *= c; }
  }
  double2* q = out + blk * 9 * Vh + idx;
#pragma unroll
  for (int e = 0; e < 9; e++) q[(size_t)e * Vh] = m.e[e];
}

int project_fields(double2* out, const double2* in, int nfields) {
  Context& c = ctx();
  const size_t ns = (size_t)nfields * c.g.V;
  k_sitewise<true><<<(unsigned)((ns + SB - 1) / SB), SB, 0, c.stream>>>(out, in, 1.0, c.g.Vh, ns);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}
static int scale_field(double2* out, const double2* in, double s) {
  Context& c = ctx();
  const size_t ns = (size_t)c.g.V;
  k_sitewise<false><<<(unsigned)((ns + SB - 1) / SB), SB, 0, c.stream>>>(out, in, s, c.g.Vh, ns);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

// Naik link, fat7l.nim:156-167: ll(x) = naik * W(x) W(x+mu) W(x+2mu)
template <bool MR>
__global__ void __launch_bounds__(SB) k_naik(double2* ll, const double2* w, int mu, double naik, int X0, int X1,
                                              int X2, int X3, int Vh, const Peers pr) {
  const int par = blockIdx.x & 1, idx = (blockIdx.x >> 1) * SB + threadIdx.x;
  if (idx >= Vh) return;
  const int X[4] = {X0, X1, X2, X3};
  const LSite s = lsite(idx, par, X);
  const LSite s1 = lhop(s, mu, +1, X), s2 = lhop(s1, mu, +1, X);
  const M3 t = m3_mul(m3_ldx<MR>(w, s1, Vh, pr), m3_ldx<MR>(w, s2, Vh, pr));   // the reference multiplies right to left too
  M3 r = m3_mul(m3_ld(w, s, Vh), t);
#pragma unroll
  for (int e = 0; e < 9; e++) { r.e[e].x *= naik; r.e[e].y *= naik; }
  m3_st(ll, s, Vh, r);
}

// ---- plaq, gaugeUtils.nim:213-284 ------------------------------------------------------------
// pl[mu(mu-1)/2 + nu] = sum_x Re tr[(U_nu(x) U_mu(x+nu))^+ (U_mu(x) U_nu(x+mu))] / (6 * 3 * vol), nu < mu
struct PlaqArgs {
  const double2* u[4];
  double* partials;   // [grid][6]
  int X[4];
  int Vh;
  Peers pr;
};
template <bool MR>
__global__ void __launch_bounds__(SB) k_plaq(const PlaqArgs a) {
  __shared__ double sm[SB / 32][6];
  const int par = blockIdx.x & 1, idx = (blockIdx.x >> 1) * SB + threadIdx.x;
  double v[6] = {0, 0, 0, 0, 0, 0};
  if (idx < a.Vh) {
    const LSite s = lsite(idx, par, a.X);
    M3 U[4];
#pragma unroll
    for (int mu = 0; mu < 4; mu++) U[mu] = m3_ld(a.u[mu], s, a.Vh);
#pragma unroll
    for (int mu = 1; mu < 4; mu++) {
      const LSite smu = lhop(s, mu, +1, a.X);
#pragma unroll
      for (int nu = 0; nu < mu; nu++) {
        const LSite snu = lhop(s, nu, +1, a.X);
        const M3 p = m3_mul(U[nu], m3_ldx<MR>(a.u[mu], snu, a.Vh, a.pr));
        const M3 q = m3_mul(U[mu], m3_ldx<MR>(a.u[nu], smu, a.Vh, a.pr));
        double d = 0.0;
#pragma unroll
        for (int e = 0; e < 9; e++) d += q.e[e].x * p.e[e].x + q.e[e].y * p.e[e].y;
        v[(mu * (mu - 1)) / 2 + nu] = d;
      }
    }
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 6; k++) {
    double t = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
    if (lane == 0) sm[w][k] = t;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    double t = 0.0;
    for (int i = 0; i < SB / 32; i++) t += sm[i][threadIdx.x];
    a.partials[(size_t)blockIdx.x * 6 + threadIdx.x] = t;
  }
}
// fixed-order final sum: one block, thread k < 6 * 32 handles plane k / 32 with stride 32
// xr != nullptr: the six sums are then added over the ranks in rank order (xrank_sum, one warp)
__global__ void __launch_bounds__(192) k_plaq_final(const double* partials, int nblocks, double norm, double* out,
                                                     XRank* xr) {
  __shared__ double sm[6];
  const int k = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double t = 0.0;
  for (int i = lane; i < nblocks; i += 32) t += partials[(size_t)i * 6 + k];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
  if (lane == 0) sm[k] = t;
  __syncthreads();
  if (k == 0) {
    for (int j = 0; j < 6; j++) {
      double v = sm[j];
      if (xr) v = xrank_sum(xr, v);
      if (lane == 0) out[j] = v * norm;
    }
  }
}

int plaq_links(const double2* const u[4], double pl[6]) {
  Context& c = ctx();
  Tmp part;
  QB_CHECK(part.get());
  PlaqArgs a;
  for (int d = 0; d < 4; d++) { a.u[d] = u[d]; a.X[d] = c.g.X[d]; }
  a.Vh = c.g.Vh;
  a.partials = reinterpret_cast<double*>(part.p);
  QB_CHECK(mr_begin(&a.pr, {u[0], u[1], u[2], u[3]}));
  const int grid = grid_par(c.g.Vh, SB);
  const double norm = 1.0 / (18.0 * (double)c.g.V * c.g.nranks);
  if (multi()) {
    QB_CHECK(xbarrier());
    k_plaq<true><<<grid, SB, 0, c.stream>>>(a);
  } else {
    k_plaq<false><<<grid, SB, 0, c.stream>>>(a);
  }
  count_launch();
  k_plaq_final<<<1, 192, 0, c.stream>>>(a.partials, grid, norm, c.d_scal + 40, xrank());
  count_launch();
  QB_CUDA(cudaGetLastError());
  QB_CHECK(fetch_scalars(c.d_scal + 40, pl, 6));
  return xbarrier();   // peers may still be reading these links
}

// ---- makeImpLinks, fat7l.nim:78-162 ------------------------------------------------------------
// coef = {oneLink, threeStaple, fiveStaple, sevenStaple, lepage}.  fl, gf: 4 fields each; ll / gfl
// may be null when naik == 0.  fl must not alias gf.  Per (dir, nu): the 3-staple (saved), the two
// 5-staples built on it (saved), then ONE launch adding Lepage + both 7-staples.
int make_imp_links(double2* const fl[4], const double2* const gf[4], const double coef[5], double2* const ll[4],
                   const double2* const gfl[4], double naik) {
  Context& c = ctx();
  QB_CHECK(mr_begin(&g_peers, {gf[0], gf[1], gf[2], gf[3], fl[0], fl[1], fl[2], fl[3]}));
  for (int d = 0; d < 4; d++)
    if (c.g.X[d] < 4) return fail("makeImpLinks: local extent %d in direction %d is below 4", c.g.X[d], d);
  double coef1 = coef[0];
  const double coef3 = coef[1], coef5 = coef[2], coef7 = coef[3], coefL = coef[4];
  const bool have5 = coef5 != 0.0 || coef7 != 0.0 || coefL != 0.0;
  const bool have3 = coef3 != 0.0 || have5;
  coef1 -= 6.0 * coefL;   // :108
  Tmp s3, s5[2];
  for (int dir = 0; dir < 4; dir++) {
    if (!have3) { QB_CHECK(scale_field(fl[dir], gf[dir], coef1)); continue; }
    QB_CHECK(s3.get());
    bool first = true;
    for (int nu = 0; nu < 4; nu++) {
      if (nu == dir) continue;
      {   // :139 computeGenStaple(staple, dir, nu, gf[dir], coef3)
        StapleArgs a = staple_args(dir);
        a.t[0] = {gf[nu], gf[dir], nu, coef3};
        if (first) { a.base = gf[dir]; a.cbase = coef1; } else a.acc = fl[dir];
        a.out = fl[dir]; a.save = s3.p;
        QB_CHECK(launch_staple(a, 1, false));
        first = false;
      }
      StapleArgs tail = staple_args(dir);
      tail.acc = fl[dir]; tail.out = fl[dir];
      int nt = 0;
      if (coefL != 0.0) tail.t[nt++] = {gf[nu], s3.p, nu, coefL};   // :141-143
      if (coef5 != 0.0 || coef7 != 0.0) {
        int k = 0;
        for (int rho = 0; rho < 4; rho++) {
          if (rho == dir || rho == nu) continue;
          QB_CHECK(s5[k].get());
          StapleArgs a = staple_args(dir);   // :147
          a.t[0] = {gf[rho], s3.p, rho, coef5};
          a.acc = fl[dir]; a.out = fl[dir]; a.save = s5[k].p;
          QB_CHECK(launch_staple(a, 1, false));
          if (coef7 != 0.0) {
            const int sig = 6 - dir - nu - rho;   // :150-153
            tail.t[nt++] = {gf[sig], s5[k].p, sig, coef7};
          }
          k++;
        }
      }
      if (nt) QB_CHECK(launch_staple(tail, nt, false));
    }
  }
  if (naik != 0.0) {
    if (!ll || !gfl) return fail("makeImpLinks: naik != 0 needs long-link input and output");
    for (int d = 0; d < 4; d++)
      if (c.g.X[d] < 4) return fail("makeImpLinks: Naik links need local extent >= 4");
    const int grid = grid_par(c.g.Vh, SB);
    for (int dir = 0; dir < 4; dir++) {
      if (multi()) {
        if (!in_arena(gfl[dir])) return fail("makeImpLinks on a split lattice: gfLong lives outside the peer arena");
        QB_CHECK(xbarrier());
        k_naik<true><<<grid, SB, 0, c.stream>>>(ll[dir], gfl[dir], dir, naik, c.g.X[0], c.g.X[1], c.g.X[2], c.g.X[3], c.g.Vh, g_peers);
      } else {
        k_naik<false><<<grid, SB, 0, c.stream>>>(ll[dir], gfl[dir], dir, naik, c.g.X[0], c.g.X[1], c.g.X[2], c.g.X[3], c.g.Vh, g_peers);
      }
      count_launch();
    }
    QB_CUDA(cudaGetLastError());
  }
  return xbarrier();   // the peers may still be reading what the caller overwrites next
}

// ---- HisqCoefs.smear, hisqLinks.nim:32-39 -------------------------------------------------------
// c = {fat7first[5], fat7second[5], naik}
void hisq_state_free(HisqState* st) {
  if (!st->live) return;
  for (int d = 3; d >= 0; d--) { pool_put(st->w[d]); pool_put(st->v[d]); }
  *st = HisqState();
}
// st != nullptr (smearGetForce, hisqsmear.nim:51-71): keeps v (first fat7) and w (its projection)
int hisq_smear(double2* const fl[4], double2* const ll[4], const double2* const g[4], const double c[11], HisqState* st) {
  Tmp t1[4], t2[4];
  double2 *p1[4], *p2[4]; const double2* cp2[4];
  if (st) {
    if (st->live) hisq_state_free(st);
    for (int d = 0; d < 4; d++) { QB_CHECK(pool_get(&st->v[d])); QB_CHECK(pool_get(&st->w[d])); p1[d] = st->v[d]; p2[d] = st->w[d]; st->u[d] = g[d]; }
    for (int k = 0; k < 11; k++) st->c[k] = c[k];
    st->live = true;
  } else {
    for (int d = 0; d < 4; d++) { QB_CHECK(t1[d].get()); p1[d] = t1[d].p; }
  }
  QB_CHECK(make_imp_links(p1, g, c, nullptr, nullptr, 0.0));
  for (int d = 0; d < 4; d++) {
    if (!st) { QB_CHECK(t2[d].get()); p2[d] = t2[d].p; }
    QB_CHECK(project_fields(p2[d], p1[d], 1));
    if (!st) t1[d].drop();
    cp2[d] = p2[d];
  }
  return make_imp_links(fl, cp2, c + 5, ll, cp2, c[10]);
}

// ---- HypCoefs.smear / smearGetForce, hypsmear.nim:49-162 ---------------------------------------------
// L1[mu][nu] = P((1-a1) g[mu] + a1/2 S(g[nu], g[mu]));  L2[mu][nu] = P((1-a2) g[mu] + a2/4 sum_a S(L1[a][b], L1[mu][b]));
// fl[mu] = P((1-a3) g[mu] + a3/6 sum_nu S(L2[nu][mu], L2[mu][nu])).  One launch per output field.
// st != nullptr (smearGetForce): the projected and unprojected fields of every level are kept for smearedForce.
static int state_alloc(HypState* st) {
  for (int mu = 0; mu < 4; mu++) {
    QB_CHECK(pool_get(&st->flx[mu]));
    for (int nu = 0; nu < 4; nu++) {
      if (nu == mu) continue;
      QB_CHECK(pool_get(&st->l1[mu][nu])); QB_CHECK(pool_get(&st->l2[mu][nu]));
      QB_CHECK(pool_get(&st->l1x[mu][nu])); QB_CHECK(pool_get(&st->l2x[mu][nu]));
    }
  }
  st->live = true;
  return 0;
}
void hyp_state_free(HypState* st) {
  if (!st->live) return;
  for (int mu = 3; mu >= 0; mu--) {   // reverse order: the pool is LIFO
    for (int nu = 3; nu >= 0; nu--) {
      if (nu == mu) continue;
      pool_put(st->l2x[mu][nu]); pool_put(st->l1x[mu][nu]); pool_put(st->l2[mu][nu]); pool_put(st->l1[mu][nu]);
    }
    pool_put(st->flx[mu]);
  }
  *st = HypState();
}

int hyp_smear(double2* const fl[4], const double2* const g[4], const double alpha[3], HypState* st) {
  Context& c = ctx();
  QB_CHECK(mr_begin(&g_peers, {g[0], g[1], g[2], g[3], fl[0], fl[1], fl[2], fl[3]}));
  for (int d = 0; d < 4; d++)
    if (c.g.X[d] < 4) return fail("hyp smear: local extent %d in direction %d is below 4", c.g.X[d], d);
  const double alp1 = alpha[0] / 2.0, alp2 = alpha[1] / 4.0, alp3 = alpha[2] / 6.0;
  const double ma1 = 1 - alpha[0], ma2 = 1 - alpha[1], ma3 = 1 - alpha[2];
  Tmp t1[4][4], t2[4][4];
  double2 *l1[4][4] = {}, *l2[4][4] = {};
  if (st) {
    if (st->live) hyp_state_free(st);
    QB_CHECK(state_alloc(st));
    for (int d = 0; d < 3; d++) st->alpha[d] = alpha[d];
    for (int d = 0; d < 4; d++) st->g[d] = g[d];
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {
    if (nu == mu) continue;
    if (st) { l1[mu][nu] = st->l1[mu][nu]; l2[mu][nu] = st->l2[mu][nu]; }
    else { QB_CHECK(t1[mu][nu].get()); l1[mu][nu] = t1[mu][nu].p; }
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {
    if (nu == mu) continue;
    StapleArgs a = staple_args(mu);   // :107-111
    a.base = g[mu]; a.cbase = ma1;
    a.t[0] = {g[nu], g[mu], nu, alp1};
    a.out = l1[mu][nu];
    if (st) a.raw = st->l1x[mu][nu];
    QB_CHECK(launch_staple(a, 1, true));
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {
    if (nu == mu) continue;
    if (!st) { QB_CHECK(t2[mu][nu].get()); l2[mu][nu] = t2[mu][nu].p; }
    StapleArgs a = staple_args(mu);   // :117-133
    a.base = g[mu]; a.cbase = ma2;
    int nt = 0;
    for (int s = 0; s < 4; s++) {
      if (s == mu || s == nu) continue;
      const int b = 6 - mu - nu - s;
      a.t[nt++] = {l1[s][b], l1[mu][b], s, alp2};
    }
    a.out = l2[mu][nu];
    if (st) a.raw = st->l2x[mu][nu];
    QB_CHECK(launch_staple(a, nt, true));
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) t1[mu][nu].drop();
  for (int mu = 0; mu < 4; mu++) {
    StapleArgs a = staple_args(mu);   // :136-150
    a.base = g[mu]; a.cbase = ma3;
    int nt = 0;
    for (int nu = 0; nu < 4; nu++) {
      if (nu == mu) continue;
      a.t[nt++] = {l2[nu][mu], l2[mu][nu], nu, alp3};
    }
    a.out = fl[mu];
    if (st) a.raw = st->flx[mu];
    QB_CHECK(launch_staple(a, nt, true));
  }
  return xbarrier();
}

// ---- smearedForce, hypsmear.nim:163-262 ---------------------------------------------------------------
// r = projectUderiv(x, chain) (matrixFunctions.nim:321-357); keep = s_keep * r (may alias chain: the next level's
// chain, already weighted by alpha/n); f = (acc ? f : 0) + s_f * r -- the `f[mu] := ma*fc; fc *= alp` pairs of
// :188-190,217-221,253-257 folded into the projection kernel.
struct ProjDerivArgs {
  const double2* x; const double2* chain;
  double2* keep; double s_keep;
  double2* f; double s_f; int acc;
  int Vh;
};
#ifndef QB_PD_MINB
#define QB_PD_MINB 1
#endif
__global__ void __launch_bounds__(SB, QB_PD_MINB) k_proj_deriv(const ProjDerivArgs a) {
  const int t = blockIdx.x * SB + threadIdx.x;
  if (t >= 2 * a.Vh) return;
  LSite s; s.par = t / a.Vh; s.idx = t - s.par * a.Vh; s.nb = NB_HOME;
  const M3 r = project_u_deriv(m3_ld(a.x, s, a.Vh), m3_ld(a.chain, s, a.Vh));
  if (a.f) {
    M3 o = m3_zero();
    if (a.acc) o = m3_ld(a.f, s, a.Vh);
    m3_axpy(o, a.s_f, r);
    m3_st(a.f, s, a.Vh, o);
  }
  M3 k = m3_zero();
  m3_axpy(k, a.s_keep, r);
  m3_st(a.keep, s, a.Vh, k);
}
static int launch_proj_deriv(const double2* x, const double2* chain, double2* keep, double s_keep, double2* f, double s_f,
                             bool acc) {
  Context& c = ctx();
  const ProjDerivArgs a = {x, chain, keep, s_keep, f, s_f, acc ? 1 : 0, c.g.Vh};
  k_proj_deriv<<<(c.g.V + SB - 1) / SB, SB, 0, c.stream>>>(a);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

// symStapleDeriv, smearutil.nim:24-54: f1 (wrt g1, direction d1) and f2 (wrt g2, direction d2) accumulate
struct DerivArgs {
  double2 *f1, *f2;
  const double2 *g1, *g2, *c;
  int d1, d2;
  int X[4];
  int Vh;
  Peers pr;
};
#ifndef QB_DERIV_MINB
#define QB_DERIV_MINB 4
#endif
template <bool MR>
__global__ void __launch_bounds__(SB, QB_DERIV_MINB) k_staple_deriv(const DerivArgs a) {
  const int par = blockIdx.x & 1, idx = (blockIdx.x >> 1) * SB + threadIdx.x;
  if (idx >= a.Vh) return;
  const LSite s = lsite(idx, par, a.X);
  M3 f1 = m3_ld(a.f1, s, a.Vh), f2 = m3_ld(a.f2, s, a.Vh);
  staple_deriv_x<MR>(f1, f2, a.g1, a.d1, a.g2, a.d2, a.c, s, a.X, a.Vh, a.pr);
  m3_st(a.f1, s, a.Vh, f1);
  m3_st(a.f2, s, a.Vh, f2);
}
// The same through the two-slot shared-memory ring of linkops.cuh staple_deriv_pipe_thread (single rank).
// EXPERIMENTAL, off unless QEXB200_DERIV_PIPE=1: the pipeline logic is checked on the host (tests/test_links_emu.py
// runs this very body on a host ring), but the kernel has not been timed on a B200 yet (tools/ab_deriv_pipe.py).
struct SmemRing {
  double2* my;   // this thread's column of the block's ring: element (slot, m, e) at my[((slot * 3 + m) * 9 + e) * PB]
  QB_HD void put(int slot, int m, const double2* f, const LSite& s, int Vh) {
#if defined(__CUDA_ARCH__)
    const double2* p = f + (size_t)s.par * 9 * Vh + s.idx;
#pragma unroll
    for (int e = 0; e < 9; e++) cp16(my + ((slot * 3 + m) * 9 + e) * PB, p + (size_t)e * Vh);
#endif
  }
  QB_HD M3 get(int slot, int m) const {
    M3 r;
#pragma unroll
    for (int e = 0; e < 9; e++) r.e[e] = my[((slot * 3 + m) * 9 + e) * PB];
    return r;
  }
  QB_HD void commit() {
#if defined(__CUDA_ARCH__)
    cp_commit();
#endif
  }
  QB_HD void wait1() {
#if defined(__CUDA_ARCH__)
    cp_wait1();
#endif
  }
  QB_HD void wait0() {
#if defined(__CUDA_ARCH__)
    cp_wait0();
#endif
  }
};
__global__ void __launch_bounds__(PB, QB_PIPE_MINB) k_staple_deriv_pipe(const DerivArgs a, int ntiles) {
  extern __shared__ double2 ring[];
  SmemRing r = {ring + threadIdx.x};
  staple_deriv_pipe_thread(r, a.f1, a.f2, a.g1, a.d1, a.g2, a.d2, a.c, a.X, a.Vh, ntiles, (int)blockIdx.x, (int)gridDim.x,
                           (int)threadIdx.x, PB);
}
static bool deriv_pipe_enabled() {
  const char* e = getenv("QEXB200_DERIV_PIPE");
  return e && e[0] == '1';
}
static int launch_deriv_pipe(const DerivArgs& a) {
  Context& c = ctx();
  static int resident = 0;
  if (!resident) {
    QB_CUDA(cudaFuncSetAttribute(k_staple_deriv_pipe, cudaFuncAttributeMaxDynamicSharedMemorySize, PIPE_SMEM));
    QB_CUDA(cudaFuncSetAttribute(k_staple_deriv_pipe, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    int dev = 0, sms = 0, occ = 0;
    QB_CUDA(cudaGetDevice(&dev));
    QB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    QB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_staple_deriv_pipe, PB, PIPE_SMEM));
    if (occ < 1) return fail("k_staple_deriv_pipe does not fit on an SM");
    resident = sms * occ;
  }
  const int ntiles = grid_par(c.g.Vh, PB);
  int grid = resident;
  if (const char* e = getenv("QEXB200_PIPE_GRID")) grid = atoi(e) > 1 ? atoi(e) : 2;
  grid -= grid & 1;
  if (grid > ntiles) grid = ntiles;
  k_staple_deriv_pipe<<<grid, PB, PIPE_SMEM, c.stream>>>(a, ntiles);
  return 0;
}
static int launch_deriv(double2* f1, double2* f2, const double2* g1, int d1, const double2* g2, int d2, const double2* ch) {
  Context& c = ctx();
  DerivArgs a;
  a.f1 = f1; a.f2 = f2; a.g1 = g1; a.g2 = g2; a.c = ch; a.d1 = d1; a.d2 = d2; a.Vh = c.g.Vh; a.pr = g_peers;
  for (int d = 0; d < 4; d++) a.X[d] = c.g.X[d];
  const int grid = grid_par(c.g.Vh, SB);
  if (multi()) {
    if (!in_arena(g1) || !in_arena(g2) || !in_arena(ch)) return fail("smeared force on a split lattice: a field lives outside the peer arena");
    QB_CHECK(xbarrier());
    k_staple_deriv<true><<<grid, SB, 0, c.stream>>>(a);
  } else if (deriv_pipe_enabled()) {
    QB_CHECK(launch_deriv_pipe(a));
  } else {
    k_staple_deriv<false><<<grid, SB, 0, c.stream>>>(a);
  }
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

// f[mu] = d/dU of sum Re tr(chain^+ V(U)) through the three levels; f may alias chain (staghmc_spv.nim:745).
int hyp_force(const HypState& st, double2* const f[4], const double2* const chain[4]) {
  Context& c = ctx();
  if (!st.live) return fail("smeared force: the smearing state is empty");
  QB_CHECK(mr_begin(&g_peers, {st.g[0], st.g[1], st.g[2], st.g[3], f[0], f[1], f[2], f[3]}));
  const double alp1 = st.alpha[0] / 2.0, alp2 = st.alpha[1] / 4.0, alp3 = st.alpha[2] / 6.0;
  const double ma1 = 1 - st.alpha[0], ma2 = 1 - st.alpha[1], ma3 = 1 - st.alpha[2];
  const size_t bytes = sizeof(double2) * 18 * (size_t)c.g.Vh;
  Tmp fc[4], fl1[4][4], fl2[4][4];
  for (int mu = 0; mu < 4; mu++) {
    QB_CHECK(fc[mu].get());
    for (int nu = 0; nu < 4; nu++) {
      if (nu == mu) continue;
      QB_CHECK(fl1[mu][nu].get()); QB_CHECK(fl2[mu][nu].get());
      QB_CUDA(cudaMemsetAsync(fl1[mu][nu].p, 0, bytes, c.stream));     // :178-182
      QB_CUDA(cudaMemsetAsync(fl2[mu][nu].p, 0, bytes, c.stream));
    }
  }
  for (int mu = 0; mu < 4; mu++)                                            // :185-190
    QB_CHECK(launch_proj_deriv(st.flx[mu], chain[mu], fc[mu].p, alp3, f[mu], ma3, false));
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {           // :191-205
    if (nu == mu) continue;
    QB_CHECK(launch_deriv(fl2[nu][mu].p, fl2[mu][nu].p, st.l2[nu][mu], nu, st.l2[mu][nu], mu, fc[mu].p));
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {           // :209-221
    if (nu == mu) continue;
    QB_CHECK(launch_proj_deriv(st.l2x[mu][nu], fl2[mu][nu].p, fl2[mu][nu].p, alp2, f[mu], ma2, true));
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {           // :222-240
    if (nu == mu) continue;
    for (int a = 0; a < 4; a++) {
      if (a == mu || a == nu) continue;
      const int b = 6 - mu - nu - a;
      QB_CHECK(launch_deriv(fl1[a][b].p, fl1[mu][b].p, st.l1[a][b], a, st.l1[mu][b], mu, fl2[mu][nu].p));
    }
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {           // :244-257
    if (nu == mu) continue;
    QB_CHECK(launch_proj_deriv(st.l1x[mu][nu], fl1[mu][nu].p, fl1[mu][nu].p, alp1, f[mu], ma1, true));
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {           // :258-264
    if (nu == mu) continue;
    QB_CHECK(launch_deriv(f[nu], f[mu], st.g[nu], nu, st.g[mu], mu, fl1[mu][nu].p));
  }
  return xbarrier();
}

// ---- fat7lDeriv, src/gauge/fat7lderiv.nim:122-335 -----------------------------------------------------------
// Like the reference, built from symStaple (k_staple) and symStapleDeriv (k_staple_deriv); the two 7-staple
// pairs and the Lepage term of one (sig, rho) go into stpl5 in ONE launch.  deriv / llderiv accumulate.
template <bool MR>
__global__ void __launch_bounds__(SB) k_naik_deriv(double2* out, const double2* U, const double2* M, int mu, double naik,
                                                    int X0, int X1, int X2, int X3, int Vh, const Peers pr) {
  const int par = blockIdx.x & 1, idx = (blockIdx.x >> 1) * SB + threadIdx.x;
  if (idx >= Vh) return;
  const int X[4] = {X0, X1, X2, X3};
  const LSite s = lsite(idx, par, X);
  M3 o = m3_ld(out, s, Vh);
  m3_axpy(o, naik, naik_deriv_x<MR>(U, M, mu, s, X, Vh, pr));
  m3_st(out, s, Vh, o);
}
// local whole-field updates between the launches; on a split lattice the peers may still be reading the
// field from their previous launch, so the ranks are lined up first
static int field_sync() { return multi() ? xbarrier() : 0; }
static int field_zero(double2* f) {
  QB_CHECK(field_sync());
  QB_CUDA(cudaMemsetAsync(f, 0, sizeof(double2) * 18 * (size_t)ctx().g.Vh, ctx().stream));
  return 0;
}
static int field_axpy(double2* y, double a, const double2* x) { QB_CHECK(field_sync()); return blas_axpy(y, a, x, (size_t)18 * ctx().g.Vh); }
static int field_axpby(double2* y, double a, const double2* x, double b) { QB_CHECK(field_sync()); return blas_axpby(y, a, x, b, (size_t)18 * ctx().g.Vh); }

int fat7_deriv(double2* const deriv[4], const double2* const gauge[4], const double2* const mid[4], const double coef[5],
               double2* const llderiv[4], const double2* const llgauge[4], const double2* const llmid[4], double naik) {
  Context& c = ctx();
  QB_CHECK(mr_begin(&g_peers, {gauge[0], gauge[1], gauge[2], gauge[3], mid[0], mid[1], mid[2], mid[3],
                               deriv[0], deriv[1], deriv[2], deriv[3]}));
  for (int d = 0; d < 4; d++)
    if (c.g.X[d] < 4) return fail("fat7lDeriv: local extent %d in direction %d is below 4", c.g.X[d], d);
  const double coefL = coef[4], coef1 = coef[0] - 6 * coefL, coef3 = coef[1], coef5 = coef[2], coef7 = coef[3];
  const bool have5 = coef5 != 0.0 || coef7 != 0.0 || coefL != 0.0;
  const bool have3 = coef3 != 0.0 || have5;
  Tmp stpl3[4], mid5[4], stpl5, mid3;
  if (have3) { QB_CHECK(stpl5.get()); QB_CHECK(mid3.get()); }
  for (int sig = 0; sig < 4; sig++) {
    if (have5) {
      for (int mu = 0; mu < 4; mu++) {
        if (mu == sig) continue;
        QB_CHECK(stpl3[mu].get()); QB_CHECK(mid5[mu].get());
        StapleArgs a = staple_args(sig);                                    // :185-186
        a.t[0] = {gauge[mu], gauge[sig], mu, 1.0};
        a.out = stpl3[mu].p;
        QB_CHECK(launch_staple(a, 1, false));
        QB_CHECK(field_zero(mid5[mu].p));
      }
      for (int rho = 0; rho < 4; rho++) {
        if (rho == sig) continue;
        int m1 = -1, n1 = -1;                                               // the two directions besides sig, rho
        for (int d = 0; d < 4; d++) if (d != sig && d != rho) { if (m1 < 0) m1 = d; else n1 = d; }
        if (coef7 != 0.0) {                                                 // :195-212
          StapleArgs a = staple_args(sig);
          a.t[0] = {gauge[n1], stpl3[m1].p, n1, coef7};
          a.t[1] = {gauge[m1], stpl3[n1].p, m1, coef7};
          if (coefL != 0.0) { a.base = stpl3[rho].p; a.cbase = coefL; }
          a.out = stpl5.p;
          QB_CHECK(launch_staple(a, 2, false));
        } else if (coefL != 0.0) {
          QB_CHECK(field_sync());
          QB_CHECK(scale_field(stpl5.p, stpl3[rho].p, coefL));
        }
        QB_CHECK(field_zero(mid3.p));
        if (coefL != 0.0 || coef7 != 0.0)                                   // :218-220
          QB_CHECK(launch_deriv(deriv[rho], mid3.p, gauge[rho], rho, stpl5.p, sig, mid[sig]));
        if (coefL != 0.0) QB_CHECK(field_axpy(mid5[rho].p, coefL, mid3.p)); // :224
        QB_CHECK(field_axpby(mid3.p, coef5, mid[sig], coef7));              // :228-230
        QB_CHECK(launch_deriv(deriv[m1], mid5[n1].p, gauge[m1], m1, stpl3[n1].p, sig, mid3.p));   // :239-241
        QB_CHECK(launch_deriv(deriv[n1], mid5[m1].p, gauge[n1], n1, stpl3[m1].p, sig, mid3.p));
      }
    }
    if (have3) {
      QB_CHECK(field_zero(mid3.p));
      for (int mu = 0; mu < 4; mu++) {
        if (mu == sig) continue;
        QB_CHECK(field_sync());
        QB_CHECK(scale_field(stpl5.p, mid[sig], coef3));                    // :250-256
        if (have5) QB_CHECK(blas_axpy(stpl5.p, 1.0, mid5[mu].p, (size_t)18 * c.g.Vh));
        QB_CHECK(launch_deriv(deriv[mu], mid3.p, gauge[mu], mu, gauge[sig], sig, stpl5.p));   // :259-261
      }
      QB_CHECK(field_axpy(deriv[sig], 1.0, mid3.p));                        // :265
    }
    if (coef1 != 0.0) QB_CHECK(field_axpy(deriv[sig], coef1, mid[sig]));    // :269
  }
  if (naik != 0.0) {                                                        // :300-326
    if (!llderiv || !llgauge || !llmid) return fail("fat7lDeriv: naik != 0 needs the long-link fields");
    const int grid = grid_par(c.g.Vh, SB);
    for (int mu = 0; mu < 4; mu++) {
      if (multi()) {
        if (!in_arena(llgauge[mu]) || !in_arena(llmid[mu])) return fail("fat7lDeriv on a split lattice: a long-link field lives outside the peer arena");
        QB_CHECK(xbarrier());
        k_naik_deriv<true><<<grid, SB, 0, c.stream>>>(llderiv[mu], llgauge[mu], llmid[mu], mu, naik, c.g.X[0], c.g.X[1], c.g.X[2], c.g.X[3], c.g.Vh, g_peers);
      } else {
        k_naik_deriv<false><<<grid, SB, 0, c.stream>>>(llderiv[mu], llgauge[mu], llmid[mu], mu, naik, c.g.X[0], c.g.X[1], c.g.X[2], c.g.X[3], c.g.Vh, g_peers);
      }
      count_launch();
    }
    QB_CUDA(cudaGetLastError());
  }
  return xbarrier();
}

// smearedForce(dsdu, dsdsu, dsdsul), hisqsmear.nim:73-80: second fat7 (+ Naik) -> projectUderiv -> first fat7.
// dsdu accumulates (fat7lDeriv's +=).
int hisq_force(const HisqState& st, double2* const dsdu[4], const double2* const dsdsu[4], const double2* const dsdsul[4]) {
  Context& c = ctx();
  if (!st.live) return fail("HISQ smeared force: the smearing state is empty");
  Tmp f[4], fll[4];
  double2 *pf[4], *pl[4]; const double2 *cw[4], *ct[4];
  for (int d = 0; d < 4; d++) {
    QB_CHECK(f[d].get()); QB_CHECK(fll[d].get());
    QB_CHECK(field_zero(f[d].p)); QB_CHECK(field_zero(fll[d].p));
    pf[d] = f[d].p; pl[d] = fll[d].p; cw[d] = st.w[d];
  }
  QB_CHECK(fat7_deriv(pf, cw, dsdsu, st.c + 5, pl, cw, dsdsul, st.c[10]));   // asqtadDeriv :16-28
  for (int d = 0; d < 4; d++) {
    if (st.c[10] != 0.0) QB_CHECK(blas_axpy(pf[d], 1.0, pl[d], (size_t)18 * c.g.Vh));
    QB_CHECK(launch_proj_deriv(st.v[d], pf[d], pf[d], 1.0, nullptr, 0.0, false));   // :77
    ct[d] = pf[d];
  }
  return fat7_deriv(dsdu, st.u, ct, st.c, nullptr, nullptr, nullptr, 0.0);   // :78
}

// projTAH(f, g, adj) (src/stagg_pv_hmc/staghmc_spv_gforce.nim:256-291): f[mu] := projectTAH(f[mu] g[mu]^+), or of
// g[mu] f[mu]^+ with adj
__global__ void __launch_bounds__(SB) k_proj_tah(double2* f, const double2* g, int adj, int Vh, size_t nsites) {
  const size_t t = (size_t)blockIdx.x * SB + threadIdx.x;
  if (t >= nsites) return;
  const size_t blk = t / Vh;
  const int idx = (int)(t - blk * Vh);
  double2* fp = f + blk * 9 * Vh + idx;
  const double2* gp = g + blk * 9 * Vh + idx;
  M3 F, G;
#pragma unroll
  for (int e = 0; e < 9; e++) { F.e[e] = fp[(size_t)e * Vh]; G.e[e] = gp[(size_t)e * Vh]; }
  const M3 r = project_tah(adj ? m3_mul_na(G, F) : m3_mul_na(F, G));
#pragma unroll
  for (int e = 0; e < 9; e++) fp[(size_t)e * Vh] = r.e[e];
}
int proj_tah(double2* f, const double2* g, int nfields, bool adj) {
  Context& c = ctx();
  const size_t ns = (size_t)nfields * c.g.V;
  k_proj_tah<<<(unsigned)((ns + SB - 1) / SB), SB, 0, c.stream>>>(f, g, adj ? 1 : 0, c.g.Vh, ns);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace qb
