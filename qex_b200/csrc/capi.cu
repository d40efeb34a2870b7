// extern "C" entry points of libqex_b200.so (include/qex_b200.h) and the host-side
// logic that sits above the kernels: stagD/D/Ddag/eoReconstruct composition
// (src/physics/stagD.nim:405-409,566-586) and the solve drivers
// (src/physics/stagSolve.nim:57-294).
#include "../../include/qex_b200.h"
#include "common.cuh"
#include <chrono>
#include <cstdarg>
#include <cstring>
#include <map>
#include <dlfcn.h>

namespace qb {

static Context g_ctx;
static std::string g_err;
Context& ctx() { return g_ctx; }

int fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}
void clear_error() { g_err.clear(); }

static NcclApi g_nccl = {};
NcclApi& nccl() { return g_nccl; }
int nccl_load() {
  if (g_nccl.loaded) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // prefer the copy already in the process
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW);
  if (!h) return fail("cannot load libnccl.so.2: %s", dlerror());
#define QB_SYM(field, name)                                                    \
  *(void**)(&g_nccl.field) = dlsym(h, name);                                   \
  if (!g_nccl.field) return fail("libnccl: missing symbol %s", name);
  QB_SYM(GetUniqueId, "ncclGetUniqueId");
  QB_SYM(CommInitRank, "ncclCommInitRank");
  QB_SYM(CommDestroy, "ncclCommDestroy");
  QB_SYM(GroupStart, "ncclGroupStart");
  QB_SYM(GroupEnd, "ncclGroupEnd");
  QB_SYM(Send, "ncclSend");
  QB_SYM(Recv, "ncclRecv");
  QB_SYM(AllReduce, "ncclAllReduce");
  QB_SYM(GetErrorString, "ncclGetErrorString");
#undef QB_SYM
  g_nccl.loaded = true;
  return 0;
}

static std::map<int, Gauge> g_gauges;
static Gauge g_hf;   // staging link set of the host-field force calls
static std::map<int, HypState> g_hyps;
static std::map<int, HisqState> g_hisqs;
static std::map<int, Field> g_fields;
static std::map<int, RngField> g_rngs;
struct CgObj { CgKeep k; int gauge, parity, x, b; };   // a CgState bound to its operator and fields
static std::map<int, CgObj> g_cgs;
static int g_next_handle = 1;
// device copies cached against the caller's host link pointers (qexb200_gauge_cached)
struct CacheEntry { std::vector<const double*> ptrs; unsigned long long version; int handle; unsigned long long used;
                    unsigned long long print; };   // print: fingerprint of a strided sample of the host links at upload
static std::vector<CacheEntry> g_cache;
static unsigned long long g_cache_clock = 0;
constexpr size_t CACHE_MAX = 8;

static int need_ctx() {
  if (!g_ctx.live) return fail("qexb200 is not initialised (call qexb200_init; a CUDA device is required, there is no CPU fallback)");
  return 0;
}
static int get_gauge(int h, Gauge** g) {
  auto it = g_gauges.find(h);
  if (it == g_gauges.end()) return fail("invalid gauge handle %d", h);
  *g = &it->second;
  return 0;
}
static int get_field(int h, Field** f) {
  auto it = g_fields.find(h);
  if (it == g_fields.end()) return fail("invalid field handle %d", h);
  *f = &it->second;
  return 0;
}
static int new_field(Field* f) {
  QB_CUDA(cudaMalloc(&f->d, sizeof(double2) * 6 * (size_t)g_ctx.g.Vh));
  return 0;
}
static size_t nhalf() { return (size_t)3 * g_ctx.g.Vh; }

// ---- operator composition ----------------------------------------------------
// stagD2, stagD.nim:348-395
static int op_stagD2(const Gauge& g, int parity, Field& r, const Field& x, double a, double b) {
  return sweep(g, parity, hf(r, parity), hf(x, 1 - parity), hf(x, parity), a, b, 1.0, 1.0, nullptr);
}
// stagD, stagD.nim:405-409: stagD2(a/(0.5 sc), m/(0.5 sc)) then r *= 0.5 sc -- the
// rescaling sweep of the reference is folded into the kernel epilogue (`post`).
static int op_stagD(const Gauge& g, int parity, Field& r, const Field& x, double m, double sc, double a) {
  const double hs = 0.5 * sc;
  return sweep(g, parity, hf(r, parity), hf(x, 1 - parity), hf(x, parity), a / hs, m / hs, 1.0, hs, nullptr);
}
static int op_D(const Gauge& g, Field& r, const Field& x, double m) {
  QB_CHECK(op_stagD(g, 0, r, x, m, 1.0, 0.0));
  return op_stagD(g, 1, r, x, m, 1.0, 0.0);
}
static int op_Ddag(const Gauge& g, Field& r, const Field& x, double m) {
  QB_CHECK(op_stagD(g, 0, r, x, m, -1.0, 0.0));
  return op_stagD(g, 1, r, x, m, -1.0, 0.0);
}
// eoReconstruct, stagD.nim:582-586: r.odd = (b.odd - D_oe r.even)/m
static int op_eoReconstruct(const Gauge& g, Field& r, const Field& b, double m) {
  QB_CHECK(op_stagD(g, 1, r, r, 0.0, -1.0 / m, 0.0));
  return blas_axpy(hf(r, 1), 1.0 / m, hf(b, 1), nhalf());
}

// norm2 of a subset, summed over ranks, on the host
static int norm2_host(const Field& f, int subset, double* out) {
  double* S = g_ctx.d_scal + 32;
  const double2* p = (subset == QEXB200_ODD) ? hf(f, 1) : f.d;
  const size_t n = (subset == QEXB200_ALL) ? 2 * nhalf() : nhalf();
  QB_CHECK(blas_norm2(p, n, S));
  QB_CHECK(allreduce_scalars(S, 1));
  return fetch_scalars(S, out, 1);
}

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// SolverParams.sloppySolve (solverBase.nim:8-9,15).  QEX's own backend ignores it; its QUDA backend maps it to mixed
// precision (qudaWrapperImpl.nim:194-197).  This library has fp64 kernels only: the request is accepted, the solve runs in
// full fp64 (the returned residuals hold in fp64 either way) and sloppy_used reports that nothing narrower ran.
static int sloppy_ok(qexb200_solver_params* sp) {
  if (sp->sloppy_solve < 0 || sp->sloppy_solve > 2)
    return fail("solver_params: sloppy_solve = %d, must be 0 (SloppyNone), 1 (SloppySingle) or 2 (SloppyHalf)", sp->sloppy_solve);
  sp->sloppy_used = 0;
  return 0;
}

// solveXX with the device backend, stagSolve.nim:57-132
static int op_solveXX(const Gauge& g, int parity, Field& r, const Field& x, double m, qexb200_solver_params* sp) {
  QB_CHECK(sloppy_ok(sp));
  const double t0 = now_s();
  QB_CHECK(blas_zero(r.d, 2 * nhalf()));                      // threads: r := 0   (stagSolve.nim:63-64)
  CgResult res;
  QB_CHECK(cg_xx(g, parity, hf(r, parity), hf(x, parity), m * m, sp->r2req, sp->maxits, true, &res));
  sp->calls = 1;
  sp->iterations = res.iterations;
  sp->iterations_max = res.iterations;
  sp->seconds = now_s() - t0;
  long long neven = (long long)g_ctx.g.Vh * g_ctx.g.nranks;
  sp->flops = (double)(g.nlinks * 4 * 72 + 60) * (double)neven * res.iterations;  // stagSolve.nim:92
  sp->r2 = res.b2 > 0 ? res.r2 / res.b2 : 0.0;
  return 0;
}

struct SolveWork { Field r, y, d; bool ok = false; int Vh = 0; };
static SolveWork g_sw;
static int solve_work() {
  if (g_sw.ok && g_sw.Vh == g_ctx.g.Vh) return 0;
  QB_CHECK(new_field(&g_sw.r)); QB_CHECK(new_field(&g_sw.y)); QB_CHECK(new_field(&g_sw.d));
  g_sw.ok = true; g_sw.Vh = g_ctx.g.Vh;
  return 0;
}

// solveReconR, stagSolve.nim:141-176 (right-preconditioned; used when one parity of b is
// negligible or m == 0)
static int solveReconR(const Gauge& g, Field& x, const Field& b, double m, qexb200_solver_params* sp,
                       double b2e, double b2o, int* its) {
  const double b2 = b2e + b2o, r2stop = sp->r2req * b2, r2stop2 = 0.5 * r2stop;
  const double r2stope = (b2o <= r2stop2) ? r2stop - b2o : r2stop2;
  Field& y = g_sw.d;  // scratch (d is unused on this branch)
  if (b2e > r2stope) {
    qexb200_solver_params s2 = *sp;
    s2.r2req = r2stope / b2e;
    QB_CHECK(op_solveXX(g, 0, y, b, m, &s2));
    *its += s2.iterations;
    QB_CHECK(blas_scale(hf(y, 0), 4.0, nhalf()));
    return op_Ddag(g, x, y, m);
  }
  const double r2stopo = (b2e <= r2stop2) ? r2stop - b2e : r2stop2;
  if (b2o > r2stopo) {
    qexb200_solver_params s2 = *sp;
    s2.r2req = r2stopo / b2o;
    QB_CHECK(op_solveXX(g, 1, y, b, m, &s2));
    *its += s2.iterations;
    QB_CHECK(blas_scale(hf(y, 1), 4.0, nhalf()));
    return op_Ddag(g, x, y, m);
  }
  return 0;
}
// solveReconL, stagSolve.nim:179-208 (left-preconditioned with odd reconstruction)
static int solveReconL(const Gauge& g, Field& x, const Field& b, double m, qexb200_solver_params* sp,
                       double b2e, double b2o, int* its) {
  Field& d = g_sw.d;
  QB_CHECK(op_Ddag(g, d, b, m));
  double d2e;
  QB_CHECK(norm2_host(d, QEXB200_EVEN, &d2e));
  qexb200_solver_params s2 = *sp;
  s2.r2req = 0.99 * sp->r2req * (b2e + b2o) * m * m / d2e;
  QB_CHECK(op_solveXX(g, 0, x, d, m, &s2));   // also zeroes x
  *its += s2.iterations;
  QB_CHECK(blas_scale(hf(x, 0), 4.0, nhalf()));
  return op_eoReconstruct(g, x, b, m);
}
// solveInner, stagSolve.nim:211-222
static int solveInner(const Gauge& g, Field& x, const Field& b, double m, qexb200_solver_params* sp,
                      double b2e, double b2o, int* its) {
  const double b2 = b2e + b2o, r2stop = sp->r2req * b2, r2stop2 = 0.5 * r2stop;
  const double r2stope = (b2o <= r2stop2) ? r2stop - b2o : r2stop2;
  const double r2stopo = (b2e <= r2stop2) ? r2stop - b2e : r2stop2;
  if (b2e <= r2stope || b2o <= r2stopo || m == 0.0) return solveReconR(g, x, b, m, sp, b2e, b2o, its);
  return solveReconL(g, x, b, m, sp, b2e, b2o, its);
}
// solve, stagSolve.nim:224-294
static int op_solve(const Gauge& g, Field& x, const Field& b, double m, qexb200_solver_params* sp0) {
  QB_CHECK(sloppy_ok(sp0));
  const double t0 = now_s();
  QB_CHECK(solve_work());
  Field &r = g_sw.r, &y = g_sw.y;
  const size_t nfull = 2 * nhalf();
  double b2, r2e, r2o;
  QB_CHECK(norm2_host(b, QEXB200_ALL, &b2));
  const double r2stop = sp0->r2req * b2;
  if (sp0->use_prev_soln) {
    QB_CHECK(op_D(g, r, x, m));
    QB_CHECK(blas_xmy(r.d, b.d, r.d, nfull));
  } else {
    QB_CHECK(blas_zero(x.d, nfull));
    QB_CHECK(blas_copy(r.d, b.d, nfull));
  }
  QB_CHECK(norm2_host(r, QEXB200_EVEN, &r2e));
  QB_CHECK(norm2_host(r, QEXB200_ODD, &r2o));
  double r2 = r2e + r2o;
  int its = 0, itsmax = 0, calls = 0;
  QB_CHECK(blas_zero(y.d, nfull));
  while (r2 > r2stop) {
    qexb200_solver_params sp = *sp0;
    sp.maxits = sp0->maxits - its;
    if (sp.maxits <= 0) break;
    sp.r2req = r2stop / r2;
    const int before = its;
    QB_CHECK(solveInner(g, y, r, m, &sp, r2e, r2o, &its));
    if (its - before > itsmax) itsmax = its - before;
    calls++;
    QB_CHECK(blas_axpy(x.d, 1.0, y.d, nfull));
    QB_CHECK(op_D(g, r, x, m));
    QB_CHECK(blas_xmy(r.d, b.d, r.d, nfull));
    QB_CHECK(norm2_host(r, QEXB200_EVEN, &r2e));
    QB_CHECK(norm2_host(r, QEXB200_ODD, &r2o));
    r2 = r2e + r2o;
  }
  sp0->calls = 1;
  sp0->iterations = its;
  sp0->iterations_max = itsmax;
  sp0->seconds = now_s() - t0;
  const double neven = (double)g_ctx.g.Vh * g_ctx.g.nranks;
  sp0->flops = (double)(g.nlinks * 4 * 72 + 60) * neven * its + (double)(g.nlinks * 4 * 72 + 24) * neven;  // stagSolve.nim:92,285
  sp0->r2 = b2 > 0 ? r2 / b2 : 0.0;
  return 0;
}

// solveXX(s, xs, b, ms, sp0, subset) with the multi-shift CG, stagSolve.nim:296-345.  shifts[0] is
// the base mass m0 (A = stagD2xx(m0^2)), shifts[k>0] = sigma_k.  xs are zeroed on the subset only
// (cgm.nim:169).
static int op_solveXX_multi(const Gauge& g, int parity, Field* const* xs, const Field& b, const double* shifts,
                            int nmass, qexb200_solver_params* sp) {
  QB_CHECK(sloppy_ok(sp));
  const double t0 = now_s();
  std::vector<double2*> xp(nmass);
  for (int m = 0; m < nmass; m++) xp[m] = hf(*xs[m], parity);
  CgResult res;
  QB_CHECK(cgm_xx(g, parity, xp.data(), hf(b, parity), shifts[0] * shifts[0], shifts, nmass, sp->r2req, sp->maxits, &res));
  sp->calls = 1;
  sp->iterations = res.iterations;
  sp->iterations_max = res.iterations;
  sp->seconds = now_s() - t0;
  const double neven = (double)g_ctx.g.Vh * g_ctx.g.nranks;
  sp->flops = (double)(g.nlinks * 4 * 72 + 60) * neven * res.iterations;   // stagSolve.nim:334
  sp->r2 = 0.0;                                                             // sp.r2.push 0.0, :343
  return 0;
}

struct MultiWork { std::vector<Field> ys, hx; Field xt; bool ok = false; int Vh = 0; };
static MultiWork g_mw;
static int multi_work(int nmass, bool host) {
  if (g_mw.Vh != g_ctx.g.Vh) {
    for (auto& f : g_mw.ys) cudaFree(f.d);
    for (auto& f : g_mw.hx) cudaFree(f.d);
    if (g_mw.ok) cudaFree(g_mw.xt.d);
    g_mw = MultiWork();
  }
  if (!g_mw.ok) { QB_CHECK(new_field(&g_mw.xt)); g_mw.ok = true; g_mw.Vh = g_ctx.g.Vh; }
  while ((int)g_mw.ys.size() < nmass) { Field f; QB_CHECK(new_field(&f)); g_mw.ys.push_back(f); }
  while (host && (int)g_mw.hx.size() < nmass) { Field f; QB_CHECK(new_field(&f)); g_mw.hx.push_back(f); }
  return 0;
}

// solve(s, xs, b, ms, sp0), stagSolve.nim:347-446: xs[k] = Ddag(m_k) y_k with (D Ddag)(m_k) y_k = b
static int op_solve_multi(const Gauge& g, Field* const* xs, const Field& b, const double* ms, int nmass,
                          qexb200_solver_params* sp0) {
  QB_CHECK(sloppy_ok(sp0));
  const double t0 = now_s();
  QB_CHECK(solve_work());
  QB_CHECK(multi_work(nmass, false));
  Field &r = g_sw.r, &xt = g_mw.xt;
  const size_t nfull = 2 * nhalf();
  const double mass = ms[0];
  double b2, b2e, b2o;
  QB_CHECK(norm2_host(b, QEXB200_ALL, &b2));
  QB_CHECK(norm2_host(b, QEXB200_EVEN, &b2e));
  QB_CHECK(norm2_host(b, QEXB200_ODD, &b2o));
  QB_CHECK(blas_copy(r.d, b.d, nfull));
  double r2 = b2e + b2o;
  const double r2stop = sp0->r2req * b2;
  std::vector<double> shifts(nmass);
  std::vector<Field*> ys(nmass);
  for (int m = 0; m < nmass; m++) {
    shifts[m] = (m == 0) ? ms[m] : 4.0 * (ms[m] * ms[m] - mass * mass);   // :391-394
    ys[m] = &g_mw.ys[m];
    QB_CHECK(blas_zero(xs[m]->d, nfull));
    QB_CHECK(blas_zero(ys[m]->d, nfull));
  }
  int its = 0, itsmax = 0, calls = 0, subset = -1;
  while (r2 > r2stop) {
    qexb200_solver_params sp = *sp0;
    sp.maxits = sp0->maxits - its;
    if (sp.maxits <= 0) break;
    sp.r2req = r2stop;
    const double r2stop2 = 0.5 * sp.r2req;
    const double r2stope = (b2o <= r2stop2) ? sp.r2req - b2o : r2stop2;
    const double r2stopo = (b2e <= r2stop2) ? sp.r2req - b2e : r2stop2;
    if (b2e > r2stope) { sp.r2req = r2stope / b2e; subset = 0; }
    else if (b2o > r2stopo) { sp.r2req = r2stopo / b2o; subset = 1; }
    if (subset < 0) return fail("solve (multi-shift): no subset qualifies for a solve");   // reference: qexError, :323
    QB_CHECK(op_solveXX_multi(g, subset, ys.data(), r, shifts.data(), nmass, &sp));
    its += sp.iterations;
    if (sp.iterations > itsmax) itsmax = sp.iterations;
    calls++;
    for (int m = 0; m < nmass; m++) QB_CHECK(blas_axpy(hf(*xs[m], subset), 4.0, hf(*ys[m], subset), nhalf()));
    QB_CHECK(op_Ddag(g, xt, *xs[0], mass));
    QB_CHECK(op_D(g, r, xt, mass));
    QB_CHECK(blas_xmy(r.d, b.d, r.d, nfull));
    QB_CHECK(norm2_host(r, QEXB200_EVEN, &b2e));
    QB_CHECK(norm2_host(r, QEXB200_ODD, &b2o));
    r2 = b2e + b2o;
  }
  for (int m = 0; m < nmass; m++) {                                        // :431-437
    if (m != 0 || calls == 0) QB_CHECK(op_Ddag(g, xt, *xs[m], ms[m]));
    QB_CHECK(blas_copy(xs[m]->d, xt.d, nfull));
  }
  sp0->calls = 1;
  sp0->iterations = its;
  sp0->iterations_max = itsmax;
  sp0->seconds = now_s() - t0;
  const double neven = (double)g_ctx.g.Vh * g_ctx.g.nranks;
  sp0->flops = (double)(g.nlinks * 4 * 72 + 60) * neven * its + (double)(g.nlinks * 4 * 72 + 24) * neven;  // :334,444
  sp0->r2 = b2 > 0 ? r2 / b2 : 0.0;
  return 0;
}

// Local geometry of one rank; ranks are lexicographic with dimension 0 slowest (qlayout.nim:94-113).
int make_geom(const int latsize[4], const int rankgeom[4], int myrank, Geom* gp) {
  Geom& g = *gp;
  g.nranks = 1; g.V = 1; g.comm_mask = 0;
  for (int d = 0; d < 4; d++) {
    if (rankgeom[d] < 1 || latsize[d] < 1 || latsize[d] % rankgeom[d] != 0) return fail("rankgeom[%d]=%d does not divide latsize %d", d, rankgeom[d], latsize[d]);
    g.lat[d] = latsize[d]; g.G[d] = rankgeom[d]; g.X[d] = latsize[d] / rankgeom[d];
    if (g.X[d] & 1) return fail("local extent %d in direction %d must be even", g.X[d], d);
    g.nranks *= rankgeom[d]; g.V *= g.X[d];
    if (rankgeom[d] > 1) g.comm_mask |= 1 << d;
  }
  if (myrank < 0 || myrank >= g.nranks) return fail("rank %d outside rank grid of %d", myrank, g.nranks);
  g.rank = myrank; g.Vh = g.V / 2;
  int r = myrank;
  for (int d = 3; d >= 0; d--) { g.rc[d] = r % g.G[d]; r /= g.G[d]; }
  return 0;
}

}  // namespace qb

using namespace qb;

extern "C" {

const char* qexb200_last_error(void) { return g_err.c_str(); }
int qexb200_is_initialized(void) { return g_ctx.live ? 1 : 0; }

int qexb200_nccl_unique_id(void* out128) {
  QB_CHECK(nccl_load());
  ncclUniqueId id;
  QB_NCCL(nccl().GetUniqueId(&id));
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  memcpy(out128, &id, 128);
  return 0;
}

int qexb200_init(const int latsize[4], const int rankgeom[4], int myrank, int device, const void* nccl_unique_id) {
  clear_error();
  if (g_ctx.live) return fail("qexb200_init: already initialised (call qexb200_finalize first)");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail("qexb200_init: no CUDA device (%s); this backend has no CPU fallback", cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail("qexb200_init: device %d out of range (%d devices)", device, ndev);
  Geom& g = g_ctx.g;
  QB_CHECK(make_geom(latsize, rankgeom, myrank, &g));
  QB_CUDA(cudaSetDevice(device));
  g_ctx.device = device;
  QB_CUDA(cudaStreamCreateWithFlags(&g_ctx.stream, cudaStreamNonBlocking));
  QB_CUDA(cudaStreamCreateWithFlags(&g_ctx.comm_stream, cudaStreamNonBlocking));
  QB_CUDA(cudaEventCreateWithFlags(&g_ctx.ev_pack, cudaEventDisableTiming));
  QB_CUDA(cudaEventCreateWithFlags(&g_ctx.ev_comm, cudaEventDisableTiming));
  QB_CUDA(cudaEventCreate(&g_ctx.ev_t0));
  QB_CUDA(cudaEventCreate(&g_ctx.ev_t1));
  // one slot per block of the largest reducing grid (the fused multi-rank sweep: pack + all sites + boundary)
  const size_t npart = 3 * ((size_t)g.Vh / 128) + 148 * 8 + 4096;
  g_ctx.npart = npart;
  QB_CUDA(cudaMalloc(&g_ctx.d_partials, sizeof(double) * npart));
  QB_CUDA(cudaMalloc(&g_ctx.d_ticket, sizeof(unsigned int) * 4));
  QB_CUDA(cudaMemset(g_ctx.d_ticket, 0, sizeof(unsigned int) * 4));
  QB_CUDA(cudaMalloc(&g_ctx.d_scal, sizeof(double) * 64));
  QB_CUDA(cudaMemset(g_ctx.d_scal, 0, sizeof(double) * 64));
  QB_CUDA(cudaHostAlloc((void**)&g_ctx.h_scal, sizeof(double) * 64, cudaHostAllocDefault));
  QB_CUDA(cudaHostAlloc((void**)&g_ctx.h_err, sizeof(int) * 4, cudaHostAllocMapped));
  g_ctx.h_err[0] = 0;
  QB_CUDA(cudaHostGetDevicePointer((void**)&g_ctx.d_err, (void*)g_ctx.h_err, 0));
  if (g.nranks > 1 && nccl_unique_id) {
    QB_CHECK(nccl_load());
    ncclUniqueId id;
    memcpy(&id, nccl_unique_id, 128);
    QB_NCCL(nccl().CommInitRank(&g_ctx.nccl, g.nranks, id, myrank));
  }
  g_ctx.inner[0] = g_ctx.inner[1] = g_ctx.inner[2] = g_ctx.inner[3] = 1;
  g_ctx.launches = 0;
  g_ctx.live = true;
  return 0;
}

int qexb200_plan_peers(const int latsize[4], const int rankgeom[4], int myrank, unsigned char peer_of[81]) {
  clear_error();
  Geom g;
  QB_CHECK(make_geom(latsize, rankgeom, myrank, &g));
  if (g.nranks > XR_MAXR) return fail("plan_peers: at most %d ranks (one box)", XR_MAXR);
  plan_peers(g, peer_of);
  return 0;
}

int qexb200_plan_halo(const int latsize[4], const int rankgeom[4], int myrank, int nlinks,
                      qexb200_halo_dir out[4], int* ndirs, long long* total_elems) {
  clear_error();
  if (nlinks != 4 && nlinks != 8) return fail("plan_halo: nlinks must be 4 or 8");
  Geom g;
  QB_CHECK(make_geom(latsize, rankgeom, myrank, &g));
  std::vector<HaloDir> dirs;
  size_t total = 0;
  QB_CHECK(plan_halo(g, nlinks / 4, dirs, &total));
  for (size_t i = 0; i < dirs.size(); i++) {
    out[i].mu = dirs[i].mu; out[i].depth = dirs[i].depth; out[i].faceh = dirs[i].faceh;
    out[i].rank_plus = dirs[i].rank_plus; out[i].rank_minus = dirs[i].rank_minus;
    out[i].off_f = (long long)dirs[i].off_f; out[i].n_f = (long long)dirs[i].n_f;
    out[i].off_b = (long long)dirs[i].off_b; out[i].n_b = (long long)dirs[i].n_b;
  }
  *ndirs = (int)dirs.size();
  *total_elems = (long long)total;
  return 0;
}

int qexb200_plan_tiles(const int latsize[4], const int rankgeom[4], int myrank, int nlinks, int parity,
                       unsigned int* out, int* ntiles) {
  clear_error();
  if (nlinks != 4 && nlinks != 8) return fail("plan_tiles: nlinks must be 4 or 8");
  if (parity != 0 && parity != 1) return fail("plan_tiles: parity must be 0 or 1");
  Geom g;
  QB_CHECK(make_geom(latsize, rankgeom, myrank, &g));
  std::vector<unsigned int> t;
  plan_tiles(g, nlinks / 4, parity, t);
  if (out) memcpy(out, t.data(), sizeof(unsigned int) * t.size());
  if (ntiles) *ntiles = (int)t.size();
  return 0;
}
int qexb200_debug_role_times(int on, unsigned long long* out, int nwords) {
  QB_CHECK(need_ctx());
  return role_times(on, out, nwords);
}

struct HostTmp { Field a, b; bool ok = false; int Vh = 0; };
static HostTmp g_ht;

int qexb200_p2p_export(void* out64) {
  QB_CHECK(need_ctx());
  return p2p_export(out64);
}
int qexb200_p2p_attach(const void* handles) {
  QB_CHECK(need_ctx());
  return p2p_attach(handles);
}
int qexb200_p2p_active(void) { return g_ctx.live && g_ctx.p2p ? 1 : 0; }

int qexb200_finalize(void) {
  if (!g_ctx.live) return 0;
  cudaStreamSynchronize(g_ctx.stream);
  cudaStreamSynchronize(g_ctx.comm_stream);
  for (auto& kv : g_gauges) { links_free(kv.second.d, kv.second.nlinks, kv.second.arena); free_compressed(kv.second); }
  for (auto& kv : g_fields) cudaFree(kv.second.d);
  g_gauges.clear(); g_fields.clear();
  g_cache.clear();
  for (auto& kv : g_rngs) rng_destroy(&kv.second);
  g_rngs.clear();
  for (auto& kv : g_cgs) { cudaFree(kv.second.k.r); cudaFree(kv.second.k.p); }
  g_cgs.clear();
  if (g_sw.ok) { cudaFree(g_sw.r.d); cudaFree(g_sw.y.d); cudaFree(g_sw.d.d); g_sw = SolveWork(); }
  for (auto& f : g_mw.ys) cudaFree(f.d);
  for (auto& f : g_mw.hx) cudaFree(f.d);
  if (g_mw.ok) cudaFree(g_mw.xt.d);
  g_mw = MultiWork();
  if (g_ht.ok) { cudaFree(g_ht.a.d); cudaFree(g_ht.b.d); g_ht = HostTmp(); }
  for (auto& kv : g_hyps) hyp_state_free(&kv.second);
  g_hyps.clear();
  for (auto& kv : g_hisqs) hisq_state_free(&kv.second);
  g_hisqs.clear();
  cg_free();
  smear_free();
  io_free();
  if (g_hf.d) { links_free(g_hf.d, g_hf.nlinks, g_hf.arena); g_hf = Gauge(); }
  halo_free();
  p2p_free();
  if (g_ctx.h_err) cudaFreeHost(g_ctx.h_err);
  if (g_ctx.nccl) { nccl().CommDestroy(g_ctx.nccl); g_ctx.nccl = nullptr; }
  if (g_ctx.d_perm) cudaFree(g_ctx.d_perm);
  if (g_ctx.d_stage) cudaFree(g_ctx.d_stage);
  cudaFree(g_ctx.d_partials); cudaFree(g_ctx.d_ticket); cudaFree(g_ctx.d_scal);
  cudaFreeHost(g_ctx.h_scal);
  cudaEventDestroy(g_ctx.ev_pack); cudaEventDestroy(g_ctx.ev_comm);
  cudaEventDestroy(g_ctx.ev_t0); cudaEventDestroy(g_ctx.ev_t1);
  cudaStreamDestroy(g_ctx.stream); cudaStreamDestroy(g_ctx.comm_stream);
  g_ctx = Context();
  return 0;
}

int qexb200_set_host_layout(const int innergeom[4]) {
  QB_CHECK(need_ctx());
  for (int d = 0; d < 4; d++) {
    if (innergeom[d] < 1 || g_ctx.g.X[d] % innergeom[d]) return fail("set_host_layout: innergeom[%d]=%d does not divide local extent %d", d, innergeom[d], g_ctx.g.X[d]);
  }
  for (int d = 0; d < 4; d++) g_ctx.inner[d] = innergeom[d];
  g_ctx.permV = 0;  // rebuild lazily
  return ensure_perm();
}

int qexb200_local_sites(int* nsites, int* neven) {
  QB_CHECK(need_ctx());
  if (nsites) *nsites = g_ctx.g.V;
  if (neven) *neven = g_ctx.g.Vh;
  return 0;
}

// ---- gauge -------------------------------------------------------------------
int qexb200_gauge_create(int nlinks, qexb200_gauge_t* out) {
  QB_CHECK(need_ctx());
  if (nlinks != 4 && nlinks != 8) return fail("gauge_create: nlinks must be 4 (naive) or 8 (fat+Naik), got %d", nlinks);
  // the three-hop neighbour index wraps once (dslash.cu hop_site, layout.cu k_sign_bwd): x +- 3 must stay within
  // one period, i.e. every local extent >= 4 (extents are even).  QEX's own limit is len <= outerGeom
  // (layout/shiftX.nim:96-97).
  if (nlinks == 8)
    for (int d = 0; d < 4; d++)
      if (g_ctx.g.X[d] < 4)
        return fail("gauge_create: fat+Naik link sets (nlinks=8) need every local extent >= 4 (three-hop shifts); "
                    "direction %d has %d", d, g_ctx.g.X[d]);
  Gauge g;
  g.nlinks = nlinks;
  QB_CHECK(links_alloc(nlinks, &g.d, &g.arena));
  const int h = g_next_handle++;
  g_gauges[h] = g;
  *out = h;
  return 0;
}
int qexb200_gauge_load(qexb200_gauge_t h, const double* const* links) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  return upload_links(*g, links);
}
// ---- handle cache for callers that own host links ------------------------------------
static void cache_forget(int h) {
  for (size_t i = 0; i < g_cache.size(); i++)
    if (g_cache[i].handle == h) { g_cache.erase(g_cache.begin() + (long)i); return; }
}
// FNV-1a over CACHE_SAMPLES evenly spaced doubles (plus the last one) of every host link field: a few KB read per
// call.  It catches what a version counter alone cannot -- links rewritten in place without a bump, or a freed set
// whose addresses were handed to a new one -- whenever the rewrite touches the whole field (smearing, rephasing, a new
// configuration all do).  A rewrite confined to sites between the samples is NOT seen: the counter stays the contract.
constexpr size_t CACHE_SAMPLES = 1024;
static unsigned long long links_fingerprint(const double* const* links, int nlinks) {
  const size_t n = (size_t)18 * (size_t)g_ctx.g.V;
  const size_t step = n > CACHE_SAMPLES ? n / CACHE_SAMPLES : 1;
  unsigned long long h = 1469598103934665603ull;
  auto mix = [&h](double v) {
    unsigned long long b;
    memcpy(&b, &v, sizeof b);
    for (int i = 0; i < 8; i++) { h ^= (b >> (8 * i)) & 0xffull; h *= 1099511628211ull; }
  };
  for (int k = 0; k < nlinks; k++) {
    if (!links[k]) continue;                     // upload_links reports the NULL
    for (size_t i = 0; i < n; i += step) mix(links[k][i]);
    mix(links[k][n - 1]);
  }
  return h;
}
int qexb200_gauge_cached(const double* const* links, int nlinks, unsigned long long version,
                         qexb200_gauge_t* out, int* cached) {
  QB_CHECK(need_ctx());
  if (!links || !out) return fail("gauge_cached: null argument");
  if (nlinks != 4 && nlinks != 8) return fail("gauge_cached: nlinks must be 4 or 8, got %d", nlinks);
  std::vector<const double*> key(links, links + nlinks);
  const unsigned long long print = links_fingerprint(links, nlinks);
  for (CacheEntry& e : g_cache) {
    if (e.ptrs != key) continue;
    e.used = ++g_cache_clock;
    *out = e.handle;
    if (e.version == version && e.print == print) { if (cached) *cached = 1; return 0; }
    Gauge* g; QB_CHECK(get_gauge(e.handle, &g));
    QB_CHECK(upload_links(*g, links));          // same storage, new content (sticky recon request re-applied)
    e.version = version;
    e.print = print;
    if (cached) *cached = 0;
    return 0;
  }
  if (g_cache.size() >= CACHE_MAX) {            // evict the least recently used set
    size_t lru = 0;
    for (size_t i = 1; i < g_cache.size(); i++) if (g_cache[i].used < g_cache[lru].used) lru = i;
    const int victim = g_cache[lru].handle;
    g_cache.erase(g_cache.begin() + (long)lru);
    QB_CHECK(qexb200_gauge_destroy(victim));
  }
  int h = 0;
  QB_CHECK(qexb200_gauge_create(nlinks, &h));
  if (qexb200_gauge_load(h, links) != 0) { const std::string keep = g_err; qexb200_gauge_destroy(h); g_err = keep; return 1; }
  g_cache.push_back(CacheEntry{key, version, h, ++g_cache_clock, print});
  *out = h;
  if (cached) *cached = 0;
  return 0;
}

int qexb200_gauge_invalidate(const double* const* links, int nlinks) {
  QB_CHECK(need_ctx());
  if (!links) return fail("gauge_invalidate: null argument");
  if (nlinks != 4 && nlinks != 8) return fail("gauge_invalidate: nlinks must be 4 or 8, got %d", nlinks);
  const std::vector<const double*> key(links, links + nlinks);
  for (size_t i = 0; i < g_cache.size(); i++) {
    if (g_cache[i].ptrs != key) continue;
    const int victim = g_cache[i].handle;
    return qexb200_gauge_destroy(victim);       // also drops the cache entry (cache_forget)
  }
  return 0;                                      // nothing cached for these pointers: not an error
}

int qexb200_gauge_destroy(qexb200_gauge_t h) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  cudaStreamSynchronize(g_ctx.stream);
  links_free(g->d, g->nlinks, g->arena);
  free_compressed(*g);
  g_gauges.erase(h);
  cache_forget(h);
  return 0;
}
int qexb200_gauge_recon(qexb200_gauge_t h, int want, int got[2], double dev[2]) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  // two digits: both link classes; four digits: one-hop class, then three-hop class (1814 = fat 18, long 14)
  const int w0 = want >= 100 ? want / 100 : want, w1 = want >= 100 ? want % 100 : want;
  if ((w0 != 12 && w0 != 14 && w0 != 18) || (w1 != 12 && w1 != 14 && w1 != 18) || want < 0)
    return fail("gauge_recon: want must be 12, 14, 18 or a pair of them (1812, 1814, 1418 ...)");
  g->want12 = (w0 == 12 ? 1 : 0) | (w1 == 12 ? 2 : 0);
  g->want14 = (w0 == 14 ? 1 : 0) | (w1 == 14 ? 2 : 0);
  QB_CHECK(compress_links(*g));
  if (got) { got[0] = g->recon[0]; got[1] = g->recon[1]; }
  if (dev) { dev[0] = g->dev[0]; dev[1] = g->dev[1]; }
  return 0;
}
int qexb200_gauge_rephase(qexb200_gauge_t h) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  return rephase(*g);
}

// ---- device link pipeline (smear.cu) -----------------------------------------
static inline double2* link_field(Gauge* g, int k) { return g->d + (size_t)k * 18 * g_ctx.g.Vh; }
// the four one-hop fields of a set (nlinks 4: k = mu; nlinks 8: k = 2 mu, stagD.nim:552-564)
static void one_hop(Gauge* g, double2* out[4]) {
  for (int mu = 0; mu < 4; mu++) out[mu] = link_field(g, g->nlinks == 8 ? 2 * mu : mu);
}
static int links_written(Gauge* g) {
  g->loaded = true;
  return compress_links(*g);
}
static int get_gauge4(int h, Gauge** g, const char* who, bool must_be_loaded) {
  QB_CHECK(get_gauge(h, g));
  if ((*g)->nlinks != 4) return fail("%s: gauge handle %d must hold 4 link fields (has %d)", who, h, (*g)->nlinks);
  if (must_be_loaded && !(*g)->loaded) return fail("%s: gauge handle %d holds no links", who, h);
  return 0;
}

void qexb200_hisq_coefs_init(qexb200_hisq_coefs* c) {   // HisqCoefs.init + setHisqFat7, hisqLinks.nim:10-24
  const double f7lf = 0.0, naik = 1.0;
  auto set = [](qexb200_fat7l_coefs& k, double f, double n) {
    k.one_link = (1.0 + 3.0 * f + n) / 8.0; k.three_staple = -1.0 / 16.0; k.five_staple = 1.0 / 64.0;
    k.seven_staple = -1.0 / 384.0; k.lepage = -f / 16.0;
  };
  set(c->fat7first, f7lf, 0.0);
  set(c->fat7second, 2.0 - f7lf, naik);
  c->naik = -naik / 24.0;
}
int qexb200_gauge_download(qexb200_gauge_t h, double* const* links) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  return download_links(*g, links);
}
int qexb200_gauge_plaq(qexb200_gauge_t h, double pl[6]) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  if (!g->loaded) return fail("gauge_plaq: gauge holds no links");
  double2* u[4]; one_hop(g, u);
  return plaq_links(u, pl);
}
int qexb200_gauge_projectU(qexb200_gauge_t hdst, qexb200_gauge_t hsrc) {
  QB_CHECK(need_ctx());
  Gauge *d, *s;
  QB_CHECK(get_gauge(hdst, &d)); QB_CHECK(get_gauge(hsrc, &s));
  if (d->nlinks != s->nlinks) return fail("gauge_projectU: link counts differ (%d vs %d)", d->nlinks, s->nlinks);
  if (!s->loaded) return fail("gauge_projectU: source holds no links");
  QB_CHECK(project_fields(d->d, s->d, s->nlinks));
  return links_written(d);
}
int qexb200_gauge_makeImpLinks(qexb200_gauge_t hfl, qexb200_gauge_t hgf, const qexb200_fat7l_coefs* coef,
                               qexb200_gauge_t hll, qexb200_gauge_t hgfl, double naik) {
  QB_CHECK(need_ctx());
  Gauge *fl, *gf, *ll = nullptr, *gfl = nullptr;
  QB_CHECK(get_gauge4(hfl, &fl, "makeImpLinks(fl)", false));
  QB_CHECK(get_gauge4(hgf, &gf, "makeImpLinks(gf)", true));
  if (fl == gf) return fail("makeImpLinks: fl and gf must be different link sets");
  const double c[5] = {coef->one_link, coef->three_staple, coef->five_staple, coef->seven_staple, coef->lepage};
  double2 *pfl[4], *pgf[4], *pll[4], *pgl[4];
  one_hop(fl, pfl); one_hop(gf, pgf);
  if (naik != 0.0) {
    QB_CHECK(get_gauge4(hll, &ll, "makeImpLinks(ll)", false));
    QB_CHECK(get_gauge4(hgfl, &gfl, "makeImpLinks(gfLong)", true));
    if (ll == gfl || ll == gf || ll == fl) return fail("makeImpLinks: ll must be its own link set");
    one_hop(ll, pll); one_hop(gfl, pgl);
  }
  QB_CHECK(make_imp_links(pfl, pgf, c, ll ? pll : nullptr, gfl ? pgl : nullptr, naik));
  QB_CHECK(links_written(fl));
  if (ll) QB_CHECK(links_written(ll));
  return 0;
}
int qexb200_gauge_smear_hisq(qexb200_gauge_t hfl, qexb200_gauge_t hll, qexb200_gauge_t hg,
                             const qexb200_hisq_coefs* coef) {
  QB_CHECK(need_ctx());
  Gauge *fl, *ll, *g;
  QB_CHECK(get_gauge4(hfl, &fl, "smear_hisq(fl)", false));
  QB_CHECK(get_gauge4(hll, &ll, "smear_hisq(ll)", false));
  QB_CHECK(get_gauge4(hg, &g, "smear_hisq(g)", true));
  if (fl == ll) return fail("smear_hisq: fl and ll must be different link sets");
  const qexb200_fat7l_coefs &a = coef->fat7first, &b = coef->fat7second;
  const double c[11] = {a.one_link, a.three_staple, a.five_staple, a.seven_staple, a.lepage,
                        b.one_link, b.three_staple, b.five_staple, b.seven_staple, b.lepage, coef->naik};
  double2 *pfl[4], *pll[4], *pg[4];
  one_hop(fl, pfl); one_hop(ll, pll); one_hop(g, pg);
  QB_CHECK(hisq_smear(pfl, pll, pg, c));
  QB_CHECK(links_written(fl));
  return links_written(ll);
}
int qexb200_gauge_smear_hyp(qexb200_gauge_t hfl, qexb200_gauge_t hg, const qexb200_hyp_coefs* coef) {
  QB_CHECK(need_ctx());
  Gauge *fl, *g;
  QB_CHECK(get_gauge4(hfl, &fl, "smear_hyp(fl)", false));
  QB_CHECK(get_gauge4(hg, &g, "smear_hyp(g)", true));
  if (fl == g) return fail("smear_hyp: fl and g must be different link sets");
  const double al[3] = {coef->alpha1, coef->alpha2, coef->alpha3};
  double2 *pfl[4], *pg[4];
  one_hop(fl, pfl); one_hop(g, pg);
  QB_CHECK(hyp_smear(pfl, pg, al));
  return links_written(fl);
}
// fat7lDeriv(deriv, gauge, mid, coef[, llderiv, llgauge, llmid, naik]) (fat7lderiv.nim:122-335); deriv, llderiv accumulate
int qexb200_gauge_fat7lDeriv(qexb200_gauge_t hd, qexb200_gauge_t hg, qexb200_gauge_t hm, const qexb200_fat7l_coefs* coef,
                             qexb200_gauge_t hld, qexb200_gauge_t hlg, qexb200_gauge_t hlm, double naik) {
  QB_CHECK(need_ctx());
  Gauge *d, *g, *m, *ld = nullptr, *lg = nullptr, *lm = nullptr;
  QB_CHECK(get_gauge4(hd, &d, "fat7lDeriv(deriv)", true));
  QB_CHECK(get_gauge4(hg, &g, "fat7lDeriv(gauge)", true));
  QB_CHECK(get_gauge4(hm, &m, "fat7lDeriv(mid)", true));
  if (d == g || d == m) return fail("fat7lDeriv: deriv must be its own link set");
  const double c[5] = {coef->one_link, coef->three_staple, coef->five_staple, coef->seven_staple, coef->lepage};
  double2 *pd[4], *pg[4], *pm[4], *pld[4], *plg[4], *plm[4];
  one_hop(d, pd); one_hop(g, pg); one_hop(m, pm);
  if (naik != 0.0) {
    QB_CHECK(get_gauge4(hld, &ld, "fat7lDeriv(llderiv)", true));
    QB_CHECK(get_gauge4(hlg, &lg, "fat7lDeriv(llgauge)", true));
    QB_CHECK(get_gauge4(hlm, &lm, "fat7lDeriv(llmid)", true));
    if (ld == lg || ld == lm) return fail("fat7lDeriv: llderiv must be its own link set");
    one_hop(ld, pld); one_hop(lg, plg); one_hop(lm, plm);
  }
  return fat7_deriv(pd, pg, pm, c, ld ? pld : nullptr, lg ? plg : nullptr, lm ? plm : nullptr, naik);
}
// HisqCoefs.smearGetForce (hisqsmear.nim:51-84)
int qexb200_hisq_smear_get_force(qexb200_gauge_t hfl, qexb200_gauge_t hll, qexb200_gauge_t hg,
                                 const qexb200_hisq_coefs* coef, qexb200_hisq_t* out) {
  QB_CHECK(need_ctx());
  Gauge *fl, *ll, *g;
  QB_CHECK(get_gauge4(hfl, &fl, "smearGetForce(su)", false));
  QB_CHECK(get_gauge4(hll, &ll, "smearGetForce(sul)", false));
  QB_CHECK(get_gauge4(hg, &g, "smearGetForce(u)", true));
  if (fl == ll || fl == g || ll == g) return fail("smearGetForce: u, su and sul must be different link sets");
  const qexb200_fat7l_coefs &a = coef->fat7first, &b = coef->fat7second;
  const double c[11] = {a.one_link, a.three_staple, a.five_staple, a.seven_staple, a.lepage,
                        b.one_link, b.three_staple, b.five_staple, b.seven_staple, b.lepage, coef->naik};
  double2 *pfl[4], *pll[4], *pg[4];
  one_hop(fl, pfl); one_hop(ll, pll); one_hop(g, pg);
  HisqState st;
  const int rc = hisq_smear(pfl, pll, pg, c, &st);
  if (rc) { hisq_state_free(&st); return rc; }
  st.gauge_handle = hg;
  QB_CHECK(links_written(fl));
  QB_CHECK(links_written(ll));
  const int h = g_next_handle++;
  g_hisqs[h] = st;
  *out = h;
  return 0;
}
int qexb200_hisq_smeared_force(qexb200_hisq_t hs, qexb200_gauge_t hdu, qexb200_gauge_t hdsu, qexb200_gauge_t hdsul) {
  QB_CHECK(need_ctx());
  auto it = g_hisqs.find(hs);
  if (it == g_hisqs.end()) return fail("invalid smearing-state handle %d", hs);
  Gauge *du, *dsu, *dsul, *g;
  QB_CHECK(get_gauge4(hdu, &du, "smearedForce(dsdu)", true));
  QB_CHECK(get_gauge4(hdsu, &dsu, "smearedForce(dsdsu)", true));
  QB_CHECK(get_gauge4(hdsul, &dsul, "smearedForce(dsdsul)", true));
  if (du == dsu || du == dsul) return fail("smearedForce: dsdu must be its own link set");
  if (get_gauge(it->second.gauge_handle, &g) != 0 || link_field(g, 0) != it->second.u[0])
    return fail("smearedForce: the thin links the state was built from are gone");
  double2 *p0[4], *p1[4], *p2[4];
  one_hop(du, p0); one_hop(dsu, p1); one_hop(dsul, p2);
  return hisq_force(it->second, p0, p1, p2);
}
int qexb200_hisq_destroy(qexb200_hisq_t hs) {
  QB_CHECK(need_ctx());
  auto it = g_hisqs.find(hs);
  if (it == g_hisqs.end()) return fail("invalid smearing-state handle %d", hs);
  cudaStreamSynchronize(g_ctx.stream);
  hisq_state_free(&it->second);
  g_hisqs.erase(it);
  return 0;
}
// smearGetForce (hypsmear.nim:49-162): smear AND keep what smearedForce needs
int qexb200_hyp_smear_get_force(qexb200_gauge_t hfl, qexb200_gauge_t hg, const qexb200_hyp_coefs* coef, qexb200_hyp_t* out) {
  QB_CHECK(need_ctx());
  Gauge *fl, *g;
  QB_CHECK(get_gauge4(hfl, &fl, "smearGetForce(fl)", false));
  QB_CHECK(get_gauge4(hg, &g, "smearGetForce(g)", true));
  if (fl == g) return fail("smearGetForce: fl and g must be different link sets");
  const double al[3] = {coef->alpha1, coef->alpha2, coef->alpha3};
  double2 *pfl[4], *pg[4];
  one_hop(fl, pfl); one_hop(g, pg);
  HypState st;
  const int rc = hyp_smear(pfl, pg, al, &st);
  if (rc) { hyp_state_free(&st); return rc; }
  st.gauge_handle = hg;
  QB_CHECK(links_written(fl));
  const int h = g_next_handle++;
  g_hyps[h] = st;
  *out = h;
  return 0;
}
// smearedForce(f, chain) (hypsmear.nim:163-262); f may be the same handle as chain (staghmc_spv.nim:745)
int qexb200_hyp_smeared_force(qexb200_hyp_t hs, qexb200_gauge_t hf, qexb200_gauge_t hchain) {
  QB_CHECK(need_ctx());
  auto it = g_hyps.find(hs);
  if (it == g_hyps.end()) return fail("invalid smearing-state handle %d", hs);
  Gauge *f, *ch, *g;
  QB_CHECK(get_gauge4(hf, &f, "smearedForce(f)", false));
  QB_CHECK(get_gauge4(hchain, &ch, "smearedForce(chain)", true));
  if (get_gauge(it->second.gauge_handle, &g) != 0 || link_field(g, 0) != it->second.g[0])
    return fail("smearedForce: the thin links the state was built from are gone");
  double2 *pf[4], *pc[4];
  one_hop(f, pf); one_hop(ch, pc);
  QB_CHECK(hyp_force(it->second, pf, pc));
  f->loaded = true;
  return 0;
}
int qexb200_hyp_destroy(qexb200_hyp_t hs) {
  QB_CHECK(need_ctx());
  auto it = g_hyps.find(hs);
  if (it == g_hyps.end()) return fail("invalid smearing-state handle %d", hs);
  cudaStreamSynchronize(g_ctx.stream);
  hyp_state_free(&it->second);
  g_hyps.erase(it);
  return 0;
}
// projTAH(f, g, adj) (stagg_pv_hmc/staghmc_spv_gforce.nim:256-291)
int qexb200_gauge_projTAH(qexb200_gauge_t hf, qexb200_gauge_t hg, int adj) {
  QB_CHECK(need_ctx());
  Gauge *f, *g;
  QB_CHECK(get_gauge(hf, &f)); QB_CHECK(get_gauge(hg, &g));
  if (f->nlinks != g->nlinks) return fail("projTAH: link counts differ (%d vs %d)", f->nlinks, g->nlinks);
  if (!f->loaded || !g->loaded) return fail("projTAH: a link set holds nothing");
  return proj_tah(f->d, g->d, f->nlinks, adj != 0);
}
// newStag3(fl, ll) (stagD.nim:552-564): g[2 mu] = fl[mu], g[2 mu + 1] = ll[mu]
int qexb200_gauge_join3(qexb200_gauge_t hout, qexb200_gauge_t hfl, qexb200_gauge_t hll) {
  QB_CHECK(need_ctx());
  Gauge *o, *fl, *ll;
  QB_CHECK(get_gauge(hout, &o));
  if (o->nlinks != 8) return fail("gauge_join3: output must hold 8 link fields");
  QB_CHECK(get_gauge4(hfl, &fl, "gauge_join3(fl)", true));
  QB_CHECK(get_gauge4(hll, &ll, "gauge_join3(ll)", true));
  const size_t n = (size_t)18 * g_ctx.g.Vh;
  for (int mu = 0; mu < 4; mu++) {
    QB_CUDA(cudaMemcpyAsync(link_field(o, 2 * mu), link_field(fl, mu), sizeof(double2) * n, cudaMemcpyDeviceToDevice, g_ctx.stream));
    QB_CUDA(cudaMemcpyAsync(link_field(o, 2 * mu + 1), link_field(ll, mu), sizeof(double2) * n, cudaMemcpyDeviceToDevice, g_ctx.stream));
  }
  return links_written(o);
}

// ---- gauge import (io.cu) ------------------------------------------
int qexb200_gauge_load_sitemajor(qexb200_gauge_t h, const void* data, int precision_bytes, int big_endian) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  if (!data) return fail("gauge_load_sitemajor: data is NULL");
  return import_sitemajor(*g, data, precision_bytes, big_endian);
}
int qexb200_gauge_read_scidac(qexb200_gauge_t h, const char* filename, unsigned int sums[4]) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  if (!filename) return fail("gauge_read_scidac: filename is NULL");
  return read_scidac(*g, filename, sums);
}
int qexb200_scidac_lattice(const char* filename, int dims[4]) {
  clear_error();
  if (!filename) return fail("scidac_lattice: filename is NULL");
  return scidac_lattice(filename, dims);
}

// the inverse of join3: fl[mu] = in[2 mu], ll[mu] = in[2 mu + 1]
int qexb200_gauge_split3(qexb200_gauge_t hfl, qexb200_gauge_t hll, qexb200_gauge_t hin) {
  QB_CHECK(need_ctx());
  Gauge *in, *fl, *ll;
  QB_CHECK(get_gauge(hin, &in));
  if (in->nlinks != 8 || !in->loaded) return fail("gauge_split3: the input must hold 8 loaded link fields");
  QB_CHECK(get_gauge4(hfl, &fl, "gauge_split3(fl)", false));
  QB_CHECK(get_gauge4(hll, &ll, "gauge_split3(ll)", false));
  if (fl == ll) return fail("gauge_split3: fl and ll must be different link sets");
  const size_t n = (size_t)18 * g_ctx.g.Vh;
  for (int mu = 0; mu < 4; mu++) {
    QB_CUDA(cudaMemcpyAsync(link_field(fl, mu), link_field(in, 2 * mu), sizeof(double2) * n, cudaMemcpyDeviceToDevice, g_ctx.stream));
    QB_CUDA(cudaMemcpyAsync(link_field(ll, mu), link_field(in, 2 * mu + 1), sizeof(double2) * n, cudaMemcpyDeviceToDevice, g_ctx.stream));
  }
  QB_CHECK(links_written(fl));
  return links_written(ll);
}

// ---- per-site random number field on the device (rng.cu) -----------------------
int qexb200_rng_create(unsigned long long seed, qexb200_rng_t* out) {
  QB_CHECK(need_ctx());
  RngField r;
  QB_CHECK(rng_create((unsigned int)seed, &r));
  const int h = g_next_handle++;
  g_rngs[h] = r;
  *out = h;
  return 0;
}
int qexb200_rng_destroy(qexb200_rng_t h) {
  QB_CHECK(need_ctx());
  auto it = g_rngs.find(h);
  if (it == g_rngs.end()) return fail("invalid rng handle %d", h);
  cudaStreamSynchronize(g_ctx.stream);
  rng_destroy(&it->second);
  g_rngs.erase(it);
  return 0;
}
static int get_rng(int h, RngField** r) {
  auto it = g_rngs.find(h);
  if (it == g_rngs.end()) return fail("invalid rng handle %d", h);
  *r = &it->second;
  return 0;
}
int qexb200_gauge_random(qexb200_gauge_t hg, qexb200_rng_t hr) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(hg, &g));
  RngField* r; QB_CHECK(get_rng(hr, &r));
  return rng_matrix_fields(*r, *g, 0);
}
int qexb200_gauge_randomTAH(qexb200_gauge_t hg, qexb200_rng_t hr) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(hg, &g));
  RngField* r; QB_CHECK(get_rng(hr, &r));
  return rng_matrix_fields(*r, *g, 1);
}
int qexb200_gauge_unit(qexb200_gauge_t hg) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(hg, &g));
  return unit_links(*g);
}
int qexb200_gauge_reunit(qexb200_gauge_t hg, double dev[4]) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(hg, &g));
  return reunit_links(*g, dev);
}
int qexb200_field_gaussian(qexb200_field_t hf_, qexb200_rng_t hr) {
  QB_CHECK(need_ctx());
  Field* f; QB_CHECK(get_field(hf_, &f));
  RngField* r; QB_CHECK(get_rng(hr, &r));
  return rng_gaussian_vector(*r, *f);
}

// ---- device fields -----------------------------------------------------------
int qexb200_field_create(qexb200_field_t* out) {
  QB_CHECK(need_ctx());
  Field f;
  QB_CHECK(new_field(&f));
  QB_CUDA(cudaMemsetAsync(f.d, 0, sizeof(double2) * 6 * (size_t)g_ctx.g.Vh, g_ctx.stream));
  const int h = g_next_handle++;
  g_fields[h] = f;
  *out = h;
  return 0;
}
int qexb200_field_destroy(qexb200_field_t h) {
  QB_CHECK(need_ctx());
  Field* f; QB_CHECK(get_field(h, &f));
  cudaStreamSynchronize(g_ctx.stream);
  cudaFree(f->d);
  g_fields.erase(h);
  return 0;
}
int qexb200_field_upload(qexb200_field_t h, const double* host) {
  QB_CHECK(need_ctx());
  Field* f; QB_CHECK(get_field(h, &f));
  QB_CHECK(upload_vector(*f, host));
  QB_CUDA(cudaStreamSynchronize(g_ctx.stream));
  return 0;
}
int qexb200_field_download(qexb200_field_t h, double* host) {
  QB_CHECK(need_ctx());
  Field* f; QB_CHECK(get_field(h, &f));
  return download_vector(*f, host);
}
int qexb200_field_zero(qexb200_field_t h) {
  QB_CHECK(need_ctx());
  Field* f; QB_CHECK(get_field(h, &f));
  return blas_zero(f->d, 2 * nhalf());
}
int qexb200_field_norm2(qexb200_field_t h, int subset, double* out) {
  QB_CHECK(need_ctx());
  Field* f; QB_CHECK(get_field(h, &f));
  return norm2_host(*f, subset, out);
}
int qexb200_field_redot(qexb200_field_t hf, qexb200_field_t hg, int subset, double* out) {
  QB_CHECK(need_ctx());
  Field *f, *g2; QB_CHECK(get_field(hf, &f)); QB_CHECK(get_field(hg, &g2));
  double* S = g_ctx.d_scal + 32;
  const size_t off = (subset == QEXB200_ODD) ? nhalf() : 0;
  const size_t n = (subset == QEXB200_ALL) ? 2 * nhalf() : nhalf();
  QB_CHECK(blas_redot(f->d + off, g2->d + off, n, S));
  QB_CHECK(allreduce_scalars(S, 1));
  return fetch_scalars(S, out, 1);
}

#define GET_G_F2(hg, h1, h2)                       \
  QB_CHECK(need_ctx());                            \
  Gauge* g; QB_CHECK(get_gauge(hg, &g));           \
  Field *f1, *f2; QB_CHECK(get_field(h1, &f1)); QB_CHECK(get_field(h2, &f2));

int qexb200_field_axpby(qexb200_field_t hy, double a, qexb200_field_t hx, double b, int subset) {
  QB_CHECK(need_ctx());
  Field *y, *x; QB_CHECK(get_field(hy, &y)); QB_CHECK(get_field(hx, &x));
  if (subset < QEXB200_EVEN || subset > QEXB200_ALL) return fail("field_axpby: subset must be EVEN, ODD or ALL");
  const size_t off = (subset == QEXB200_ODD) ? nhalf() : 0;
  const size_t n = (subset == QEXB200_ALL) ? 2 * nhalf() : nhalf();
  if (b == 0.0 && a == 1.0) return y == x ? 0 : blas_copy(y->d + off, x->d + off, n);
  return blas_axpby(y->d + off, a, x->d + off, b, n);
}
int qexb200_dev_stagD2(qexb200_gauge_t hg, int parity, qexb200_field_t r, qexb200_field_t x, double a, double b) {
  GET_G_F2(hg, r, x);
  if (parity != 0 && parity != 1) return fail("stagD2: parity must be 0 or 1");
  return op_stagD2(*g, parity, *f1, *f2, a, b);
}
int qexb200_dev_stagD(qexb200_gauge_t hg, int parity, qexb200_field_t r, qexb200_field_t x, double m, double sc, double a) {
  GET_G_F2(hg, r, x);
  if (parity != 0 && parity != 1) return fail("stagD: parity must be 0 or 1");
  return op_stagD(*g, parity, *f1, *f2, m, sc, a);
}
int qexb200_dev_D(qexb200_gauge_t hg, qexb200_field_t r, qexb200_field_t x, double m) {
  GET_G_F2(hg, r, x);
  return op_D(*g, *f1, *f2, m);
}
int qexb200_dev_Ddag(qexb200_gauge_t hg, qexb200_field_t r, qexb200_field_t x, double m) {
  GET_G_F2(hg, r, x);
  return op_Ddag(*g, *f1, *f2, m);
}
int qexb200_dev_stagD2xx(qexb200_gauge_t hg, int parity, qexb200_field_t r, qexb200_field_t x, double m2) {
  GET_G_F2(hg, r, x);
  if (parity != 0 && parity != 1) return fail("stagD2xx: parity must be 0 or 1");
  return apply_xx_public(*g, parity, hf(*f1, parity), hf(*f2, parity), m2);
}
int qexb200_dev_solveXX(qexb200_gauge_t hg, int parity, qexb200_field_t r, qexb200_field_t x, double m, qexb200_solver_params* sp) {
  GET_G_F2(hg, r, x);
  if (parity != 0 && parity != 1) return fail("solveXX: parity must be 0 or 1");
  return op_solveXX(*g, parity, *f1, *f2, m, sp);
}
int qexb200_dev_solve(qexb200_gauge_t hg, qexb200_field_t x, qexb200_field_t b, double m, qexb200_solver_params* sp) {
  GET_G_F2(hg, x, b);
  return op_solve(*g, *f1, *f2, m, sp);
}

// solveEO, stagSolve.nim:17-38: cgSolve(r, x, stagD2ee(m^2)) on the even subset -- r is NOT zeroed, its content is
// the initial guess (CgState.solve forms r0 = b - A x on its first call, cg.nim:146-149)
static int op_solveEO(const Gauge& g, Field& r, const Field& x, double m, qexb200_solver_params* sp) {
  QB_CHECK(sloppy_ok(sp));
  const double t0 = now_s();
  CgResult res;
  QB_CHECK(cg_xx(g, 0, hf(r, 0), hf(x, 0), m * m, sp->r2req, sp->maxits, false, &res));
  sp->calls = 1;
  sp->iterations = res.iterations;
  sp->iterations_max = res.iterations;
  sp->seconds = now_s() - t0;
  sp->flops = (double)(g.nlinks * 4 * 72 + 60) * (double)g_ctx.g.Vh * g_ctx.g.nranks * res.iterations;   // :31
  sp->r2 = res.b2 > 0 ? res.r2 / res.b2 : 0.0;
  return 0;
}
int qexb200_dev_solveEO(qexb200_gauge_t hg, qexb200_field_t r, qexb200_field_t x, double m, qexb200_solver_params* sp) {
  GET_G_F2(hg, r, x);
  return op_solveEO(*g, *f1, *f2, m, sp);
}

// ---- CgState as an object (cg.nim:13-53, 256-283): newCgState / reset / solve (restartable) -------------------
int qexb200_cg_create(qexb200_gauge_t hg, int parity, qexb200_field_t x, qexb200_field_t b, qexb200_cg_t* out) {
  GET_G_F2(hg, x, b);
  (void)g; (void)f1; (void)f2;
  if (parity != 0 && parity != 1) return fail("cg_create: parity must be 0 or 1");
  CgObj o;
  o.gauge = hg; o.parity = parity; o.x = x; o.b = b;
  QB_CUDA(cudaMalloc(&o.k.r, sizeof(double2) * nhalf()));
  QB_CUDA(cudaMalloc(&o.k.p, sizeof(double2) * nhalf()));
  QB_CUDA(cudaMemsetAsync(o.k.p, 0, sizeof(double2) * nhalf(), g_ctx.stream));
  const int h = g_next_handle++;
  g_cgs[h] = o;
  *out = h;
  return 0;
}
static int get_cg(int h, CgObj** o) {
  auto it = g_cgs.find(h);
  if (it == g_cgs.end()) return fail("invalid CG state handle %d", h);
  *o = &it->second;
  return 0;
}
int qexb200_cg_reset(qexb200_cg_t h) {     // CgState.reset, cg.nim:21-27
  QB_CHECK(need_ctx());
  CgObj* o; QB_CHECK(get_cg(h, &o));
  o->k.b2 = -1.0; o->k.iterations = 0; o->k.r2 = 1.0; o->k.r2old = 1.0;
  return 0;
}
int qexb200_cg_solve(qexb200_cg_t h, double m, qexb200_solver_params* sp) {   // CgState.solve(op, sp), cg.nim:55-254
  QB_CHECK(sloppy_ok(sp));
  QB_CHECK(need_ctx());
  CgObj* o; QB_CHECK(get_cg(h, &o));
  Gauge* g; QB_CHECK(get_gauge(o->gauge, &g));
  Field *x, *b; QB_CHECK(get_field(o->x, &x)); QB_CHECK(get_field(o->b, &b));
  const double t0 = now_s();
  const int before = o->k.b2 >= 0.0 ? o->k.iterations : 0;
  CgResult res;
  QB_CHECK(cg_xx(*g, o->parity, hf(*x, o->parity), hf(*b, o->parity), m * m, sp->r2req, sp->maxits, false, &res, &o->k));
  sp->calls = 1;
  sp->iterations = res.iterations;          // sp.finalIterations = state.iterations: cumulative (cg.nim:252)
  sp->iterations_max = res.iterations;
  sp->seconds = now_s() - t0;
  sp->flops = (double)(g->nlinks * 4 * 72 + 60) * (double)g_ctx.g.Vh * g_ctx.g.nranks * (res.iterations - before);
  sp->r2 = res.b2 > 0 ? res.r2 / res.b2 : 0.0;
  return 0;
}
int qexb200_cg_destroy(qexb200_cg_t h) {
  QB_CHECK(need_ctx());
  CgObj* o; QB_CHECK(get_cg(h, &o));
  cudaStreamSynchronize(g_ctx.stream);
  cudaFree(o->k.r); cudaFree(o->k.p);
  g_cgs.erase(h);
  return 0;
}

// ---- fermion-force outer products (force.cu) ----------------------
static int op_outer(Gauge& f, const Field& a, const Field& b, double ce, double co, bool acc) {
  const OuterPair p = {a.d, b.d, ce, co};
  QB_CHECK(outer_products(f, &p, 1, acc));
  f.loaded = true;
  // the links changed in place: refresh a 12/14-real copy if one was asked for (none: no sync, nothing to do)
  return (f.want12 || f.want14) ? compress_links(f) : 0;
}
// stagDeriv(s, f, x), stagD.nim:635-664: f[mu] += x (x) x(+mu)^+ on even sites, -= on odd sites, then s.rephase f
static int op_stagDeriv(Gauge& f, const Field& x) {
  if (f.nlinks != 4) return fail("stagDeriv: the force field must hold 4 link fields (s.rephase f, stagD.nim:664)");
  QB_CHECK(op_outer(f, x, x, 1.0, -1.0, true));
  return rephase(f);
}
// stagD2deriv(s, g, c, x), stagD.nim:588-633: g[mu] += c (x) x(+h)^+ - x (x) c(+h)^+
static int op_stagD2deriv(Gauge& g, const Field& c, const Field& x) {
  const OuterPair p[2] = {{c.d, x.d, 1.0, 1.0}, {x.d, c.d, -1.0, -1.0}};
  QB_CHECK(outer_products(g, p, 2, true));
  g.loaded = true;
  return (g.want12 || g.want14) ? compress_links(g) : 0;
}
static int same_links(const Gauge* s, const Gauge* f, const char* who) {
  if (s->nlinks != f->nlinks) return fail("%s: operator has %d link fields, force field %d", who, s->nlinks, f->nlinks);
  return 0;
}
int qexb200_gauge_zero(qexb200_gauge_t h) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  QB_CUDA(cudaMemsetAsync(g->d, 0, sizeof(double2) * (size_t)g->nlinks * 18 * g_ctx.g.Vh, g_ctx.stream));
  g->loaded = true;
  free_compressed(*g);
  return 0;
}
int qexb200_gauge_scale(qexb200_gauge_t h, double c) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(h, &g));
  if (!g->loaded) return fail("gauge_scale: gauge holds no links");
  QB_CHECK(blas_scale(g->d, c, (size_t)g->nlinks * 18 * g_ctx.g.Vh));
  return compress_links(*g);
}
int qexb200_dev_outer(qexb200_gauge_t hf, qexb200_field_t ha, qexb200_field_t hb, double scale_even,
                      double scale_odd, int accumulate) {
  QB_CHECK(need_ctx());
  Gauge* f; QB_CHECK(get_gauge(hf, &f));
  Field *a, *b; QB_CHECK(get_field(ha, &a)); QB_CHECK(get_field(hb, &b));
  if (accumulate && !f->loaded) return fail("outer: accumulating into a link set that holds nothing");
  return op_outer(*f, *a, *b, scale_even, scale_odd, accumulate != 0);
}
int qexb200_dev_stagDeriv(qexb200_gauge_t hs, qexb200_gauge_t hf, qexb200_field_t hx) {
  QB_CHECK(need_ctx());
  Gauge *s, *f; QB_CHECK(get_gauge(hs, &s)); QB_CHECK(get_gauge(hf, &f));
  Field* x; QB_CHECK(get_field(hx, &x));
  QB_CHECK(same_links(s, f, "stagDeriv"));
  if (!f->loaded) return fail("stagDeriv: f holds nothing (zero it first, the reference accumulates)");
  return op_stagDeriv(*f, *x);
}
int qexb200_dev_stagD2deriv(qexb200_gauge_t hs, qexb200_gauge_t hg, qexb200_field_t hc, qexb200_field_t hx) {
  QB_CHECK(need_ctx());
  Gauge *s, *g; QB_CHECK(get_gauge(hs, &s)); QB_CHECK(get_gauge(hg, &g));
  Field *c, *x; QB_CHECK(get_field(hc, &c)); QB_CHECK(get_field(hx, &x));
  QB_CHECK(same_links(s, g, "stagD2deriv"));
  if (!g->loaded) return fail("stagD2deriv: g holds nothing (zero it first, the reference accumulates)");
  return op_stagD2deriv(*g, *c, *x);
}

// ---- host-buffer entry points: H2D, device op, D2H ---------------------------
static int host_tmp() {
  if (g_ht.ok && g_ht.Vh == g_ctx.g.Vh) return 0;
  if (g_ht.ok) { cudaFree(g_ht.a.d); cudaFree(g_ht.b.d); g_ht = HostTmp(); }
  QB_CHECK(new_field(&g_ht.a)); QB_CHECK(new_field(&g_ht.b));
  g_ht.ok = true; g_ht.Vh = g_ctx.g.Vh;
  return 0;
}
#define HOST_PROLOGUE(hg)                    \
  QB_CHECK(need_ctx());                      \
  Gauge* g; QB_CHECK(get_gauge(hg, &g));     \
  QB_CHECK(host_tmp());                      \
  Field &R = g_ht.a, &X = g_ht.b;

int qexb200_stagD2(qexb200_gauge_t hg, int parity, double* r, const double* x, double a, double b) {
  HOST_PROLOGUE(hg);
  if (parity != 0 && parity != 1) return fail("stagD2: parity must be 0 or 1");
  if (x == r) { QB_CHECK(upload_vector(R, r)); QB_CHECK(op_stagD2(*g, parity, R, R, a, b)); return download_vector(R, r); }
  QB_CHECK(upload_vector(R, r));
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_stagD2(*g, parity, R, X, a, b));
  return download_vector(R, r);
}
int qexb200_stagD(qexb200_gauge_t hg, int parity, double* r, const double* x, double m, double sc, double a) {
  HOST_PROLOGUE(hg);
  if (parity != 0 && parity != 1) return fail("stagD: parity must be 0 or 1");
  if (x == r) { QB_CHECK(upload_vector(R, r)); QB_CHECK(op_stagD(*g, parity, R, R, m, sc, a)); return download_vector(R, r); }
  QB_CHECK(upload_vector(R, r));
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_stagD(*g, parity, R, X, m, sc, a));
  return download_vector(R, r);
}
int qexb200_D(qexb200_gauge_t hg, double* r, const double* x, double m) {
  HOST_PROLOGUE(hg);
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_D(*g, R, X, m));
  return download_vector(R, r);
}
int qexb200_Ddag(qexb200_gauge_t hg, double* r, const double* x, double m) {
  HOST_PROLOGUE(hg);
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_Ddag(*g, R, X, m));
  return download_vector(R, r);
}
// The remaining flavours of stagD.nim:405-427,572-580, each a parameter choice of the same sweep:
// peqDdag: r += Ddag x; eoReduce: r.even = (Ddag b).even; stagDb: r = m x + sc D x without the final rescaling pass
// (i.e. stagD2 with b = m/(0.5 sc)); stagD1 / stagD1x: + / - the bare hop sum.
int qexb200_peqDdag(qexb200_gauge_t hg, double* r, const double* x, double m) {
  HOST_PROLOGUE(hg);
  QB_CHECK(upload_vector(R, r));
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_stagD(*g, 0, R, X, m, -1.0, 1.0));
  QB_CHECK(op_stagD(*g, 1, R, X, m, -1.0, 1.0));
  return download_vector(R, r);
}
int qexb200_eoReduce(qexb200_gauge_t hg, double* r, const double* b, double m) {
  HOST_PROLOGUE(hg);
  QB_CHECK(upload_vector(R, r));
  QB_CHECK(upload_vector(X, b));
  QB_CHECK(op_stagD(*g, 0, R, X, m, -1.0, 0.0));
  return download_vector(R, r);
}
int qexb200_stagDb(qexb200_gauge_t hg, int parity, double* r, const double* x, double m, double sc) {
  HOST_PROLOGUE(hg);
  if (parity != 0 && parity != 1) return fail("stagDb: parity must be 0 or 1");
  if (sc == 0.0) return fail("stagDb: sc must not be 0");
  QB_CHECK(upload_vector(R, r));
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_stagD2(*g, parity, R, X, 0.0, m / (0.5 * sc)));
  return download_vector(R, r);
}
int qexb200_stagD1(qexb200_gauge_t hg, int parity, double* r, const double* x, int minus) {
  HOST_PROLOGUE(hg);
  if (parity != 0 && parity != 1) return fail("stagD1: parity must be 0 or 1");
  QB_CHECK(upload_vector(R, r));
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_stagD2(*g, parity, R, X, 0.0, 0.0));
  if (minus) QB_CHECK(blas_scale(hf(R, parity), -1.0, nhalf()));
  return download_vector(R, r);
}
int qexb200_dev_peqDdag(qexb200_gauge_t hg, qexb200_field_t r, qexb200_field_t x, double m) {
  GET_G_F2(hg, r, x);
  QB_CHECK(op_stagD(*g, 0, *f1, *f2, m, -1.0, 1.0));
  return op_stagD(*g, 1, *f1, *f2, m, -1.0, 1.0);
}
int qexb200_dev_eoReduce(qexb200_gauge_t hg, qexb200_field_t r, qexb200_field_t b, double m) {
  GET_G_F2(hg, r, b);
  return op_stagD(*g, 0, *f1, *f2, m, -1.0, 0.0);
}
int qexb200_stagD2xx(qexb200_gauge_t hg, int parity, double* r, const double* x, double m2) {
  HOST_PROLOGUE(hg);
  if (parity != 0 && parity != 1) return fail("stagD2xx: parity must be 0 or 1");
  QB_CHECK(upload_vector(R, r));   // the other parity of r is preserved
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(apply_xx_public(*g, parity, hf(R, parity), hf(X, parity), m2));
  return download_vector(R, r);
}
int qexb200_eoReconstruct(qexb200_gauge_t hg, double* r, const double* b, double m) {
  HOST_PROLOGUE(hg);
  QB_CHECK(upload_vector(R, r));
  QB_CHECK(upload_vector(X, b));
  QB_CHECK(op_eoReconstruct(*g, R, X, m));
  return download_vector(R, r);
}
// stagDeriv / stagD2deriv on HOST fields: f (nd link fields, accumulated into) and the vectors in native QEX memory
static int host_force_tmp(int nlinks) {
  if (g_hf.d && g_hf.nlinks == nlinks) return 0;
  if (g_hf.d) { links_free(g_hf.d, g_hf.nlinks, g_hf.arena); g_hf = Gauge(); }
  g_hf.nlinks = nlinks;
  return links_alloc(nlinks, &g_hf.d, &g_hf.arena);
}
int qexb200_stagDeriv(qexb200_gauge_t hg, double* const* f, const double* x) {
  HOST_PROLOGUE(hg);
  (void)R;
  QB_CHECK(host_force_tmp(g->nlinks));
  QB_CHECK(upload_links(g_hf, f));
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_stagDeriv(g_hf, X));
  return download_links(g_hf, f);
}
int qexb200_stagD2deriv(qexb200_gauge_t hg, double* const* gr, const double* c, const double* x) {
  HOST_PROLOGUE(hg);
  QB_CHECK(host_force_tmp(g->nlinks));
  QB_CHECK(upload_links(g_hf, gr));
  QB_CHECK(upload_vector(R, c));
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_stagD2deriv(g_hf, R, X));
  return download_links(g_hf, gr);
}
int qexb200_solveXX(qexb200_gauge_t hg, int parity, double* r, const double* x, double m, qexb200_solver_params* sp) {
  HOST_PROLOGUE(hg);
  if (parity != 0 && parity != 1) return fail("solveXX: parity must be 0 or 1");
  const double t0 = now_s();
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_solveXX(*g, parity, R, X, m, sp));
  QB_CHECK(download_vector(R, r));
  sp->seconds = now_s() - t0;
  return 0;
}
int qexb200_solveEO(qexb200_gauge_t hg, double* r, const double* x, double m, qexb200_solver_params* sp) {
  HOST_PROLOGUE(hg);
  const double t0 = now_s();
  QB_CHECK(upload_vector(R, r));     // the initial guess
  QB_CHECK(upload_vector(X, x));
  QB_CHECK(op_solveEO(*g, R, X, m, sp));
  QB_CHECK(download_vector(R, r));
  sp->seconds = now_s() - t0;
  return 0;
}
int qexb200_solve(qexb200_gauge_t hg, double* x, const double* b, double m, qexb200_solver_params* sp) {
  HOST_PROLOGUE(hg);
  const double t0 = now_s();
  if (sp->use_prev_soln) QB_CHECK(upload_vector(R, x));
  QB_CHECK(upload_vector(X, b));
  QB_CHECK(op_solve(*g, R, X, m, sp));
  QB_CHECK(download_vector(R, x));
  sp->seconds = now_s() - t0;
  return 0;
}

// ---- multi-shift solves -----------------------------------------------------------------------
static int get_fields(const qexb200_field_t* hs, int n, std::vector<Field*>& out) {
  out.resize(n);
  for (int i = 0; i < n; i++) QB_CHECK(get_field(hs[i], &out[i]));
  return 0;
}
int qexb200_dev_solveXX_multi(qexb200_gauge_t hg, int parity, const qexb200_field_t* xs, qexb200_field_t b,
                              const double* shifts, int nmass, qexb200_solver_params* sp) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(hg, &g));
  if (parity != 0 && parity != 1) return fail("solveXX (multi-shift): parity must be 0 or 1");
  if (nmass < 1) return fail("solveXX (multi-shift): nmass must be >= 1");
  std::vector<Field*> fx; QB_CHECK(get_fields(xs, nmass, fx));
  Field* fb; QB_CHECK(get_field(b, &fb));
  return op_solveXX_multi(*g, parity, fx.data(), *fb, shifts, nmass, sp);
}
int qexb200_dev_solve_multi(qexb200_gauge_t hg, const qexb200_field_t* xs, qexb200_field_t b, const double* ms,
                            int nmass, qexb200_solver_params* sp) {
  QB_CHECK(need_ctx());
  Gauge* g; QB_CHECK(get_gauge(hg, &g));
  if (nmass < 1) return fail("solve (multi-shift): nmass must be >= 1");
  std::vector<Field*> fx; QB_CHECK(get_fields(xs, nmass, fx));
  Field* fb; QB_CHECK(get_field(b, &fb));
  return op_solve_multi(*g, fx.data(), *fb, ms, nmass, sp);
}
int qexb200_solveXX_multi(qexb200_gauge_t hg, int parity, double* const* xs, const double* b, const double* shifts,
                          int nmass, qexb200_solver_params* sp) {
  HOST_PROLOGUE(hg);
  (void)R;
  if (parity != 0 && parity != 1) return fail("solveXX (multi-shift): parity must be 0 or 1");
  if (nmass < 1) return fail("solveXX (multi-shift): nmass must be >= 1");
  const double t0 = now_s();
  QB_CHECK(multi_work(nmass, true));
  std::vector<Field*> fx(nmass);
  for (int m = 0; m < nmass; m++) { fx[m] = &g_mw.hx[m]; QB_CHECK(upload_vector(*fx[m], xs[m])); }   // other parity is preserved
  QB_CHECK(upload_vector(X, b));
  QB_CHECK(op_solveXX_multi(*g, parity, fx.data(), X, shifts, nmass, sp));
  for (int m = 0; m < nmass; m++) QB_CHECK(download_vector(*fx[m], xs[m]));
  sp->seconds = now_s() - t0;
  return 0;
}
int qexb200_solve_multi(qexb200_gauge_t hg, double* const* xs, const double* b, const double* ms, int nmass,
                        qexb200_solver_params* sp) {
  HOST_PROLOGUE(hg);
  (void)R;
  if (nmass < 1) return fail("solve (multi-shift): nmass must be >= 1");
  const double t0 = now_s();
  QB_CHECK(multi_work(nmass, true));
  std::vector<Field*> fx(nmass);
  for (int m = 0; m < nmass; m++) fx[m] = &g_mw.hx[m];
  QB_CHECK(upload_vector(X, b));
  QB_CHECK(op_solve_multi(*g, fx.data(), X, ms, nmass, sp));
  for (int m = 0; m < nmass; m++) QB_CHECK(download_vector(*fx[m], xs[m]));
  sp->seconds = now_s() - t0;
  return 0;
}

// ---- measurement -------------------------------------------------------------
int qexb200_sync(void) {
  QB_CHECK(need_ctx());
  QB_CUDA(cudaStreamSynchronize(g_ctx.stream));
  QB_CUDA(cudaStreamSynchronize(g_ctx.comm_stream));
  return check_device_error();
}
int qexb200_timer_start(void) {
  QB_CHECK(need_ctx());
  QB_CUDA(cudaEventRecord(g_ctx.ev_t0, g_ctx.stream));
  return 0;
}
int qexb200_timer_stop(float* ms) {
  QB_CHECK(need_ctx());
  QB_CUDA(cudaEventRecord(g_ctx.ev_t1, g_ctx.stream));
  QB_CUDA(cudaEventSynchronize(g_ctx.ev_t1));
  QB_CUDA(cudaEventElapsedTime(ms, g_ctx.ev_t0, g_ctx.ev_t1));
  return 0;
}
long long qexb200_launch_count(int reset) {
  const long long n = g_ctx.launches;
  if (reset) g_ctx.launches = 0;
  return n;
}

}  // extern "C"
