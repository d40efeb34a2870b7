// Internal declarations shared by the translation units of libqex_b200.so.
// Device data layout:
//   colour-vector field  double2 v[2 parity][3 colour][Vh]          (site fastest)
//   link set             double2 u[nlinks][2 parity][9 (row,col)][Vh]
// with Vh = local sites / 2 and the checkerboard index idx = lex(x)/2,
// lex = ((x3*X2+x2)*X1+x1)*X0+x0 over LOCAL coordinates.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>
#include "sweep_site.cuh"   // Links, SweepArgs, the per-site functions (includes linkops.cuh)
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

namespace qb {

struct Geom {
  int X[4];      // local extents (all even)
  int G[4];      // rank grid
  int rc[4];     // my rank coordinates
  int lat[4];    // global extents
  int V, Vh;
  int comm_mask; // bit mu set: direction mu is split over ranks
  int nranks, rank;
};

// Halo geometry of one split direction.
struct HaloDir {
  int mu;
  int faceh;      // sites of one parity in one face layer
  int depth;      // 1 (naive) or 3 (Naik)
  // element (double2) offsets inside the per-sweep send/recv buffers
  size_t off_f, off_b;     // psi layers / product layers
  size_t n_f, n_b;         // number of double2 in each
  int rank_plus, rank_minus;
};

// ---- peer-memory window (NVLink direct stores between the ranks of one box) --------------
// Every rank owns one cudaMalloc'ed window with the same layout and exports it with
// cudaIpcGetMemHandle; peers map it (qexb200_p2p_export / qexb200_p2p_attach).
//   [0, 256)                 halo flags   unsigned f[2 buf][4 dir][2 side], written by the neighbours
//   [256, 256 + LL bytes)    LL slots of the cross-rank scalar sums: u64 ll[NB][MAXR][2]
//   [WIN_HDR, ...)           2 receive buffers of halo_cap double2 each
constexpr int XR_MAXR = 16;      // ranks an in-kernel cross-rank sum supports (one box has 8 GPUs)
constexpr int XR_NB = 4;         // LL slots in flight
constexpr size_t WIN_FLAGS = 256;
constexpr size_t WIN_LL = sizeof(unsigned long long) * XR_NB * XR_MAXR * 2;
constexpr size_t WIN_HDR = 4096;
static_assert(WIN_FLAGS + WIN_LL <= WIN_HDR, "window header overflow");

// Cross-rank scalar sum descriptor, in device memory.  ll[j] is rank j's LL array (peer mapping;
// ll[rank] is the local one).  Protocol: NCCL-LL style 8-byte words {seq:32 | half of the double:32}.
struct XRank {
  unsigned long long* ll[XR_MAXR];
  int nranks, rank;
  unsigned int seq;        // reductions done so far (device side counter)
  int* err;                // mapped host word: != 0 after a wait timed out
  unsigned long long timeout_ns;
  unsigned long long* stamps;   // diagnostic: [0] ns spent inside xrank_sum, [1] calls, [2] longest (or nullptr)
};

// Programmatic dependent launch (sm_90+): a kernel launched with pdl = true may become resident while the kernel before
// it in the stream is still draining its last wave.  Contract of every kernel that can be launched so: it calls
// pdl_wait() before it touches ANY memory an earlier kernel of the stream writes (the wait returns once the previous
// grid has completed and its writes are visible), and pdl_trigger() only AFTER its own pdl_wait() -- so kernel K+1 can
// overlap kernel K only, never K-1, and what K+1 does before its wait need only be independent of K.  Without the launch
// attribute both calls are no-ops.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

struct Context {
  bool live = false;
  bool pdl_loop = false;   // set by the CG loop on one GPU: its sweeps and BLAS-1 kernels are launched with the PDL attribute
  int device = 0;
  Geom g;
  cudaStream_t stream = nullptr, comm_stream = nullptr;
  cudaEvent_t ev_pack = nullptr, ev_comm = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
  ncclComm_t nccl = nullptr;
  int inner[4] = {1, 1, 1, 1};
  int* d_perm = nullptr;       // device site (parity*Vh+idx) -> host single-site index
  int permV = 0;               // lanes of the host layout the table was built for
  double* d_stage = nullptr;   // raw host-layout staging on the device
  size_t stage_bytes = 0;
  double* h_pinned = nullptr;  // pinned staging
  size_t pinned_bytes = 0;
  // reductions
  double* d_partials = nullptr;
  size_t npart = 0;            // slots in d_partials
  unsigned int* d_ticket = nullptr;
  double* d_scal = nullptr;    // device scalars
  double* h_scal = nullptr;    // pinned mirror
  // halo
  std::vector<HaloDir> halo;
  double2 *sendbuf = nullptr, *recvbuf = nullptr;
  size_t halo_elems = 0;
  long long launches = 0;
  // peer-memory path
  bool p2p = false;                 // window attached on every rank: halos and sums go over NVLink stores
  char* win = nullptr;              // my window
  size_t win_bytes = 0, halo_cap = 0;   // halo_cap: double2 per receive buffer
  char* peer_win[XR_MAXR] = {};     // mapped windows (peer_win[rank] == win)
  XRank* d_xr = nullptr;            // device copy of the descriptor
  int* h_err = nullptr;             // mapped pinned error word
  int* d_err = nullptr;             // its device alias
  unsigned int halo_seq = 0;        // parity sweeps exchanged so far (identical on all ranks)
  // symmetric arena inside the window: link-shaped fields of multi-rank contexts live here, at the same
  // offsets on every rank (SPMD call order), so kernels can read a neighbour's copy over NVLink
  size_t arena_off = 0, arena_unit = 0; // byte offset inside the window; bytes of one matrix field (18 Vh double2)
  int arena_units = 0;
  std::vector<char> arena_used;     // one flag per unit
  unsigned long long p2p_timeout_ns = 120ull * 1000000000ull;  // QEXB200_P2P_TIMEOUT_S
  // multi-rank sweep: block -> tile tables [link classes - 1][output parity] (dslash.cu tile_order)
  unsigned int* d_order[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
  int pack_blocks = 2 * 148;   // cap on the pack blocks of the multi-rank sweep (QEXB200_PACK_BLOCKS)
  // role timestamps of the multi-rank sweep (qexb200_debug_role_times): STAMP_RING launches x 8 words + 8 for xrank_sum
  unsigned long long* d_stamps = nullptr;
};
constexpr int STAMP_RING = 1024;

// NCCL is bound at run time (dlopen of libnccl.so.2) the first time a multi-rank
// context is created: a process that also hosts another NCCL user (e.g. torch) must
// share ITS copy of the library, and single-rank users need no NCCL at all.
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*);
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*GroupStart)();
  ncclResult_t (*GroupEnd)();
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  const char* (*GetErrorString)(ncclResult_t);
  bool loaded;
};
NcclApi& nccl();
int nccl_load();

Context& ctx();
inline XRank* xrank() { return ctx().p2p ? ctx().d_xr : nullptr; }
int fail(const char* fmt, ...);
void clear_error();

#define QB_CUDA(call)                                                                      \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) return qb::fail("%s:%d: %s: %s", __FILE__, __LINE__, #call,      \
                                           cudaGetErrorString(e_));                        \
  } while (0)
#define QB_NCCL(call)                                                                      \
  do {                                                                                     \
    ncclResult_t e_ = (call);                                                              \
    if (e_ != ncclSuccess) return qb::fail("%s:%d: %s: %s", __FILE__, __LINE__, #call,      \
                                           qb::nccl().GetErrorString(e_));                 \
  } while (0)
#define QB_CHECK(call)            \
  do {                            \
    int rc_ = (call);             \
    if (rc_ != 0) return rc_;     \
  } while (0)

// ---- fields -----------------------------------------------------------------
struct Field {
  double2* d = nullptr;  // [2][3][Vh]
};
struct Gauge {
  int nlinks = 0;        // 4 or 8
  double2* d = nullptr;  // [nlinks][2][9][Vh]
  bool loaded = false;
  bool arena = false;    // d lives in the peer-mapped arena (multi-rank contexts)
  // 12-real compression (layout.cu compress_links)
  int want12 = 0;                 // sticky request (bit j: class j), re-applied after every load / rephase
  int recon[2] = {18, 18};        // what the sweeps read per link class
  double2* d12[2] = {nullptr, nullptr};
  unsigned short* sgn = nullptr;
  double inv[2] = {1.0, 1.0};
  double dev[2] = {0.0, 0.0};     // measured max |U - rebuilt U| / |s| per class (diagnostic)
  int want14 = 0;                 // same, 14-real storage (rows 0, 1 + one complex factor per link)
  int rc() const {
    return (recon[0] == 12 ? 1 : 0) | (recon[1] == 12 ? 2 : 0) | (recon[0] == 14 ? 4 : 0) | (recon[1] == 14 ? 8 : 0);
  }
  Links links() const {
    Links L;
    L.u = d; L.u12[0] = d12[0]; L.u12[1] = d12[1]; L.sgn = sgn; L.inv[0] = inv[0]; L.inv[1] = inv[1];
    return L;
  }
};
inline double2* hf(const Field& f, int parity) { return f.d + (size_t)parity * 3 * ctx().g.Vh; }

// ---- dslash.cu --------------------------------------------------------------
// r[p] = post*(a*r + b*bx + sgn*sum_k[U x(+h) - U^+ x(-h)]); dot != nullptr also returns
// sum Re<bx, r_new> over ranks' local sites into *dot (device pointer, not yet all-reduced).
int sweep(const Gauge& g, int parity, double2* r, const double2* x, const double2* bx, double a,
          double b, double sgn, double post, double* dot_out, const double* skip = nullptr);

// ---- blas.cu ----------------------------------------------------------------
// All vectors are half fields (3*Vh double2) unless n is given.
int blas_zero(double2* x, size_t n2);
int blas_copy(double2* y, const double2* x, size_t n2);
int blas_scale(double2* x, double s, size_t n2);
int blas_axpy(double2* y, double a, const double2* x, size_t n2);      // y += a x
int blas_xpay(double2* y, const double2* x, double a, size_t n2);      // y = x + a y
int blas_xmy(double2* z, const double2* x, const double2* y, size_t n2);  // z = x - y
int blas_axpby(double2* y, double a, const double2* x, double b, size_t n2);  // y = a x + b y
// local (this rank) reductions into device scalar slot; all-reduced by reduce_finish
int blas_norm2(const double2* x, size_t n2, double* d_out);
int blas_redot(const double2* x, const double2* y, size_t n2, double* d_out);
// sum over ranks (NCCL) when nranks > 1, in place on device scalars
int allreduce_scalars(double* d_ptr, int n);
// copy n device scalars to host (synchronising)
int fetch_scalars(const double* d_ptr, double* h_out, int n);
// non-zero status if a device-side wait on a peer timed out
int check_device_error();

// ---- solver.cu --------------------------------------------------------------
struct CgResult { int iterations; double r2, b2; };
// what a CgState keeps between calls of solve (cg.nim:13-27): r, p and the scalars; b2 < 0 = after reset
struct CgKeep {
  double2 *r = nullptr, *p = nullptr;   // half fields
  double b2 = -1.0, r2 = 1.0, r2old = 1.0;
  int iterations = 0;
};
int cg_xx(const Gauge& g, int parity, double2* x, const double2* b, double m2, double r2req,
          int maxits, bool x_is_zero, CgResult* res, CgKeep* keep = nullptr);
int cgm_xx(const Gauge& g, int parity, double2* const* xs, const double2* b, double m2, const double* sigmas,
           int nmass, double r2req, int maxits, CgResult* res);
int apply_xx_public(const Gauge& g, int parity, double2* out, const double2* x, double m2);
int cg_free();

// ---- layout.cu --------------------------------------------------------------
int ensure_perm();
int upload_vector(Field& f, const double* host);
int download_vector(const Field& f, double* host);
int upload_links(Gauge& g, const double* const* links);
int download_links(const Gauge& g, double* const* links);
int rephase(Gauge& g);
int compress_links(Gauge& g);      // honours g.want12; sets g.recon
void free_compressed(Gauge& g);

// ---- smear.cu (device link pipeline; matrix field = double2[2][9][Vh]) -------
int make_imp_links(double2* const fl[4], const double2* const gf[4], const double coef[5], double2* const ll[4],
                   const double2* const gfl[4], double naik);
// what HisqCoefs.smearGetForce keeps (hisqsmear.nim:60-71): v = first fat7, w = projectU(v); u by reference
struct HisqState {
  double2 *v[4] = {}, *w[4] = {};
  const double2* u[4] = {};
  double c[11] = {};
  int gauge_handle = -1;
  bool live = false;
};
int hisq_smear(double2* const fl[4], double2* const ll[4], const double2* const g[4], const double c[11], HisqState* st = nullptr);
int hisq_force(const HisqState& st, double2* const dsdu[4], const double2* const dsdsu[4], const double2* const dsdsul[4]);
void hisq_state_free(HisqState* st);
int fat7_deriv(double2* const deriv[4], const double2* const gauge[4], const double2* const mid[4], const double coef[5],
               double2* const llderiv[4], const double2* const llgauge[4], const double2* const llmid[4], double naik);
// what smearGetForce keeps for smearedForce (hypsmear.nim:62-75): projected and unprojected fields of each level
struct HypState {
  double2 *l1[4][4] = {}, *l2[4][4] = {}, *l1x[4][4] = {}, *l2x[4][4] = {}, *flx[4] = {};
  const double2* g[4] = {};    // the thin links the state was built from (must stay unchanged, hypsmear.nim:51-53)
  double alpha[3] = {0, 0, 0};
  int gauge_handle = -1;
  bool live = false;
};
int hyp_smear(double2* const fl[4], const double2* const g[4], const double alpha[3], HypState* st = nullptr);
int hyp_force(const HypState& st, double2* const f[4], const double2* const chain[4]);
void hyp_state_free(HypState* st);
int proj_tah(double2* f, const double2* g, int nfields, bool adj);
int project_fields(double2* out, const double2* in, int nfields);
int plaq_links(const double2* const u[4], double pl[6]);
int smear_free();

// ---- force.cu (fermion-force outer products) ---------------------------------
struct OuterPair {
  const double2* a;   // full field [2][3][Vh], taken at x
  const double2* b;   // full field, taken at x + h
  double ce, co;      // weight on even / odd sites x
};
int outer_products(Gauge& f, const OuterPair* pairs, int np, bool accumulate);

// ---- io.cu (SciDAC/LIME and site-major import into the device layout) -------
int import_sitemajor(Gauge& g, const void* data, int prec, int big_endian);
int read_scidac(Gauge& g, const char* fn, unsigned sums[4]);
int scidac_lattice(const char* fn, int dims[4]);
int io_free();

// ---- rng.cu (per-site RngMilc6 field on the device: the reference's input generator) ----
struct RngField {
  unsigned int* st = nullptr;   // [9][V]: r0..r6, ic, mult of RngMilc6; site = parity * Vh + idx
};
int rng_create(unsigned int seed, RngField* r);
void rng_destroy(RngField* r);
int rng_gaussian_vector(RngField& r, Field& f);
int rng_matrix_fields(RngField& r, Gauge& g, int mode);   // 0: g.random r, 1: g.randomTAH r
int unit_links(Gauge& g);
int reunit_links(Gauge& g, double out[4]);   // checkSU, projectSU, checkSU (staghmc_sh.nim:247-257)

// ---- comm.cu ----------------------------------------------------------------
int halo_setup();
int plan_halo(const Geom& g, int nl, std::vector<HaloDir>& dirs, size_t* total);
int make_geom(const int latsize[4], const int rankgeom[4], int myrank, Geom* g);
int halo_free();
int p2p_alloc();                       // allocate the window (multi-rank contexts)
int p2p_export(void* out64);
int p2p_attach(const void* handles);   // nranks * 64 bytes, rank order
int p2p_free();
// matrix-field storage: the peer-mapped arena on multi-rank contexts with a window, cudaMalloc otherwise
int links_alloc(int nfields, double2** out, bool* in_arena);
void links_free(double2* p, int nfields, bool in_arena);
bool in_arena(const void* p);
void plan_peers(const Geom& g, unsigned char peer_of[81]);
int peers_fill(Peers* pr);          // kernel-side view of the arena (needs qexb200_p2p_attach)
int xbarrier();                     // all ranks' earlier work on the compute stream is complete and visible
int halo_begin(const Gauge& g);
void plan_tiles(const Geom& g, int nl, int parity, std::vector<unsigned int>& out);   // dslash.cu
int role_times(int on, unsigned long long* out, int nwords);                          // comm.cu
// Start the exchange for a sweep that outputs `parity`: pack, then send/recv on comm_stream.
int halo_start(const Gauge& g, int parity, const double2* x);
// Make the compute stream wait for the exchange.
int halo_wait();

inline void count_launch(int n = 1) { ctx().launches += n; }

}  // namespace qb
