// Conjugate gradient on the even-even / odd-odd normal operator, device resident.
//
// WHAT: CgState.solve with cpNone (src/solvers/cg.nim:55-272) applied to
// A = stagD2ee/oo(m2) (src/physics/stagD.nim:433-469), as driven by solveXX
// (src/physics/stagSolve.nim:57-99).
// HOW (B200): every scalar of the recurrence (rz, rzold, pAp, alpha, beta, the
// iteration counter and the stop flag) lives in device memory; kernels read them,
// so the host never synchronises inside the loop.  Reductions are fused into the
// kernels that already stream the vectors (pAp into the second Dslash sweep, |r|^2
// into the x/r update) and are deterministic (fixed two-level order).  The host
// runs ahead by a few iterations and watches a status word the step kernel writes
// to mapped pinned memory; once the stop flag is set the update kernels are
// no-ops, so the iteration count and x are exactly those of the textbook loop.
#include "stencil.cuh"
#include <chrono>
#include <cstdlib>

namespace qb {

// device scalar slots
enum { S_R2 = 0, S_R2OLD, S_PAP, S_R2NEW, S_R2STOP, S_ITN, S_MAXITS, S_DONE, S_B2, S_R2REQ, S_SEQ, S_TMP0, S_TMP1, S_TMP2, S_TMP3, S_N };

// Status ring in mapped pinned memory: execution j of k_cg_step fills slot j % RING and
// publishes it by writing seq = j last.  The host reads the slot of a FIXED iteration
// (launched - LAG), so every rank takes the same decision at the same point and the
// number of NCCL calls stays identical across ranks.
constexpr int RING = 8;
struct CgStatus {
  volatile long long seq;
  volatile long long itn;
  volatile long long done;
  volatile double r2;
};

static CgStatus* h_status = nullptr;
static CgStatus* d_status = nullptr;  // device alias of h_status
static double2 *w_r = nullptr, *w_p = nullptr, *w_Ap = nullptr, *w_t = nullptr;  // CG work vectors
static int w_Vh = 0;

constexpr int CB = 256;

// p = r + beta p   (cg.nim:186-193; itn==0: p := r)
__global__ void __launch_bounds__(CB) k_cg_p(double2* __restrict__ p, const double2* __restrict__ r,
                                              const double* __restrict__ S, size_t n) {
  pdl_wait();
  pdl_trigger();
  if (S[S_DONE] != 0.0) return;
  const bool first = (S[S_ITN] == 0.0);
  const double beta = first ? 0.0 : S[S_R2] / S[S_R2OLD];
  for (size_t i = (size_t)blockIdx.x * CB + threadIdx.x; i < n; i += (size_t)gridDim.x * CB) {
    const double2 rv = r[i];
    double2 o = rv;
    if (!first) { const double2 pv = p[i]; o.x = fma(beta, pv.x, rv.x); o.y = fma(beta, pv.y, rv.y); }
    p[i] = o;
  }
}

// publishes the loop state to the host ring (execution j fills slot j % RING, seq written last)
__device__ __forceinline__ void cg_publish(double* S, CgStatus* st) {
  const long long seq = (long long)(S[S_SEQ] += 1.0);
  CgStatus* slot = st + (seq % RING);
  slot->itn = (long long)S[S_ITN];
  slot->done = (long long)S[S_DONE];
  slot->r2 = S[S_R2];
  __threadfence_system();
  slot->seq = seq;
}

// x += alpha p ; r -= alpha Ap ; r2new = |r|^2   (cg.nim:208-213), then -- on the thread that
// completes the reduction -- the loop control of cg.nim:174 and the status for the host.
__global__ void __launch_bounds__(CB) k_cg_xr(double2* __restrict__ x, double2* __restrict__ r,
                                               const double2* __restrict__ p, const double2* __restrict__ Ap,
                                               double* S, size_t n, double* partials, unsigned int* ticket,
                                               XRank* xr, CgStatus* st) {
  pdl_wait();
  pdl_trigger();
  if (S[S_DONE] != 0.0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) cg_publish(S, st);
    return;
  }
  const double alpha = S[S_R2] / S[S_PAP];
  double s = 0.0;
  for (size_t i = (size_t)blockIdx.x * CB + threadIdx.x; i < n; i += (size_t)gridDim.x * CB) {
    const double2 pv = p[i], av = Ap[i];
    double2 xv = x[i], rv = r[i];
    xv.x = fma(alpha, pv.x, xv.x); xv.y = fma(alpha, pv.y, xv.y);
    rv.x = fma(-alpha, av.x, rv.x); rv.y = fma(-alpha, av.y, rv.y);
    x[i] = xv; r[i] = rv;
    s = fma(rv.x, rv.x, s); s = fma(rv.y, rv.y, s);
  }
  if (grid_reduce<CB>(s, partials, ticket, S + S_R2NEW, xr)) {
    S[S_R2OLD] = S[S_R2];
    S[S_R2] = S[S_R2NEW];
    S[S_ITN] += 1.0;
    if (!(S[S_R2] > S[S_R2STOP]) || S[S_ITN] >= S[S_MAXITS]) S[S_DONE] = 1.0;
    cg_publish(S, st);
  }
}

// NCCL path: update without loop control (ncclAllReduce of r2new follows), then k_cg_step
__global__ void __launch_bounds__(CB) k_cg_xr_nostep(double2* __restrict__ x, double2* __restrict__ r,
                                                      const double2* __restrict__ p, const double2* __restrict__ Ap,
                                                      double* S, size_t n, double* partials, unsigned int* ticket) {
  if (S[S_DONE] != 0.0) return;
  const double alpha = S[S_R2] / S[S_PAP];
  double s = 0.0;
  for (size_t i = (size_t)blockIdx.x * CB + threadIdx.x; i < n; i += (size_t)gridDim.x * CB) {
    const double2 pv = p[i], av = Ap[i];
    double2 xv = x[i], rv = r[i];
    xv.x = fma(alpha, pv.x, xv.x); xv.y = fma(alpha, pv.y, xv.y);
    rv.x = fma(-alpha, av.x, rv.x); rv.y = fma(-alpha, av.y, rv.y);
    x[i] = xv; r[i] = rv;
    s = fma(rv.x, rv.x, s); s = fma(rv.y, rv.y, s);
  }
  grid_reduce<CB>(s, partials, ticket, S + S_R2NEW);
}
__global__ void k_cg_step(double* S, CgStatus* st) {
  if (S[S_DONE] == 0.0) {
    S[S_R2OLD] = S[S_R2];
    S[S_R2] = S[S_R2NEW];
    S[S_ITN] += 1.0;
    if (!(S[S_R2] > S[S_R2STOP]) || S[S_ITN] >= S[S_MAXITS]) S[S_DONE] = 1.0;
  }
  cg_publish(S, st);
}

// after the initial residual: r2stop = r2req*b2 (cg.nim:155), loop entry test (:163)
// itn0 > 0: a restarted CgState (cg.nim:158-161 picks up iterations, r2old from the state)
__global__ void k_cg_init(double* S, CgStatus* st, double r2req, double maxits, double itn0, double r2old0) {
  S[S_R2] = S[S_R2NEW];
  S[S_R2OLD] = r2old0;
  S[S_R2STOP] = r2req * S[S_B2];
  S[S_ITN] = itn0;
  S[S_MAXITS] = maxits;
  S[S_DONE] = (!(S[S_R2] > S[S_R2STOP]) || maxits <= itn0) ? 1.0 : 0.0;
  S[S_SEQ] = 0.0;
  st->itn = (long long)itn0; st->done = (long long)S[S_DONE]; st->r2 = S[S_R2];
  __threadfence_system();
  st->seq = 0;
}

static int cg_workspace() {
  Context& c = ctx();
  if (w_Vh == c.g.Vh && w_r) return 0;
  if (w_r) { cudaFree(w_r); cudaFree(w_p); cudaFree(w_Ap); cudaFree(w_t); w_r = nullptr; }
  const size_t bytes = sizeof(double2) * 3 * (size_t)c.g.Vh;
  QB_CUDA(cudaMalloc(&w_r, bytes));
  QB_CUDA(cudaMalloc(&w_p, bytes));
  QB_CUDA(cudaMalloc(&w_Ap, bytes));
  QB_CUDA(cudaMalloc(&w_t, bytes));
  w_Vh = c.g.Vh;
  if (!h_status) {
    QB_CUDA(cudaHostAlloc((void**)&h_status, sizeof(CgStatus) * RING, cudaHostAllocMapped));
    QB_CUDA(cudaHostGetDevicePointer((void**)&d_status, (void*)h_status, 0));
  }
  return 0;
}

static void cgm_free();
int cg_free() {
  cgm_free();
  if (w_r) { cudaFree(w_r); cudaFree(w_p); cudaFree(w_Ap); cudaFree(w_t); w_r = w_p = w_Ap = w_t = nullptr; w_Vh = 0; }
  if (h_status) { cudaFreeHost((void*)h_status); h_status = nullptr; d_status = nullptr; }
  return 0;
}

// A = stagD2xx (stagD.nim:433-469): t[q] = (2D) x ; out[p] = 4 m2 x - (2D) t ; optional <x, out>
// skip: device scalar that turns both sweeps into no-ops when non-zero (the CG's done flag, see SweepArgs::skip)
static int apply_xx(const Gauge& g, int parity, double2* out, const double2* x, double m2, double* dot,
                    const double* skip = nullptr) {
  QB_CHECK(sweep(g, 1 - parity, w_t, x, nullptr, 0.0, 0.0, 1.0, 1.0, nullptr, skip));
  QB_CHECK(sweep(g, parity, out, w_t, x, 0.0, 4.0 * m2, -1.0, 1.0, dot, skip));
  return 0;
}

int apply_xx_public(const Gauge& g, int parity, double2* out, const double2* x, double m2) {
  QB_CHECK(cg_workspace());
  return apply_xx(g, parity, out, x, m2, nullptr);
}

// QEXB200_PDL=0: plain stream order (A/B runs); anything else, or unset: programmatic dependent launch
static bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("QEXB200_PDL"); return !(e && e[0] == '0'); }();
  return on;
}

static inline int cg_grid(size_t n) {
  size_t gsz = (n + CB - 1) / CB;
  const size_t cap = 148 * 8;
  return (int)(gsz < cap ? (gsz ? gsz : 1) : cap);
}

int cg_xx(const Gauge& g, int parity, double2* x, const double2* b, double m2, double r2req,
          int maxits, bool x_is_zero, CgResult* res, CgKeep* keep) {
  Context& c = ctx();
  QB_CHECK(cg_workspace());
  const size_t n = (size_t)3 * c.g.Vh;
  double* S = c.d_scal;
  const int grid = cg_grid(n);
  // a CgState object keeps r and p between calls (cg.nim:13-19); one-shot solves use the shared work vectors
  double2* const w_r = keep ? keep->r : qb::w_r;
  double2* const w_p = keep ? keep->p : qb::w_p;
  const bool restart = keep && keep->b2 >= 0.0;
  const int itn0 = restart ? keep->iterations : 0;

  QB_CUDA(cudaStreamSynchronize(c.stream));
  for (int i = 0; i < RING; i++) h_status[i].seq = -1;

  if (!restart) {
    // first call of CgState.solve, cg.nim:132-153
    QB_CHECK(blas_norm2(b, n, S + S_B2));
    if (x_is_zero) {
      QB_CHECK(blas_copy(w_r, b, n));
    } else {
      QB_CHECK(apply_xx(g, parity, w_Ap, x, m2, nullptr));
      QB_CHECK(blas_xmy(w_r, b, w_Ap, n));
    }
    QB_CHECK(blas_norm2(w_r, n, S + S_R2NEW));
    QB_CHECK(allreduce_scalars(S + S_B2, 1));
    QB_CHECK(allreduce_scalars(S + S_R2NEW, 1));
  } else {
    // later call on the same state (cg.nim:155-161): b2, r2, r2old, iterations and the vectors r, p carry over
    c.h_scal[0] = keep->b2; c.h_scal[1] = keep->r2;
    QB_CUDA(cudaMemcpyAsync(S + S_B2, c.h_scal, sizeof(double), cudaMemcpyHostToDevice, c.stream));
    QB_CUDA(cudaMemcpyAsync(S + S_R2NEW, c.h_scal + 1, sizeof(double), cudaMemcpyHostToDevice, c.stream));
  }
  k_cg_init<<<1, 1, 0, c.stream>>>(S, d_status, r2req, (double)maxits, (double)itn0, restart ? keep->r2old : 1.0);
  count_launch();
  double hs[2];
  QB_CHECK(fetch_scalars(S + S_B2, hs, 1));
  const double b2 = hs[0];
  if (b2 == 0.0) {  // cg.nim:139-144
    QB_CHECK(blas_zero(x, n));
    if (keep) { QB_CHECK(blas_zero(w_r, n)); keep->b2 = 0.0; keep->r2 = 0.0; keep->r2old = 1.0; keep->iterations = 0; }
    res->iterations = 0; res->r2 = 0.0; res->b2 = 0.0;
    return 0;
  }

  const int LAG = 3;  // iterations the host may run ahead of the device
  int launched = 0;
  bool done = h_status[0].done != 0;  // slot 0 = state after k_cg_init (fetch_scalars synchronised)
  // one GPU: the four kernels of an iteration are launched with the programmatic-dependent-launch attribute, so each
  // becomes resident while its predecessor drains (common.cuh pdl_wait / pdl_trigger).  QEXB200_PDL=0 turns it off (A/B).
  const bool pdl = (c.g.nranks == 1) && pdl_enabled();
  c.pdl_loop = pdl;
  struct PdlOff { Context& c; ~PdlOff() { c.pdl_loop = false; } } pdl_off{c};
  while (!done && launched < maxits - itn0) {
    QB_CUDA(launch_k(k_cg_p, dim3(grid), dim3(CB), c.stream, pdl, w_p, w_r, S, n));
    count_launch();
    QB_CHECK(apply_xx(g, parity, w_Ap, w_p, m2, S + S_PAP, S + S_DONE));
    QB_CHECK(allreduce_scalars(S + S_PAP, 1));
    if (c.g.nranks > 1 && !c.p2p) {
      // NCCL path: the sum over ranks sits between the update and the loop control
      k_cg_xr_nostep<<<grid, CB, 0, c.stream>>>(x, w_r, w_p, w_Ap, S, n, c.d_partials, c.d_ticket);
      count_launch();
      QB_CHECK(allreduce_scalars(S + S_R2NEW, 1));
      k_cg_step<<<1, 1, 0, c.stream>>>(S, d_status);
      count_launch();
    } else {
      QB_CUDA(launch_k(k_cg_xr, dim3(grid), dim3(CB), c.stream, pdl, x, w_r, w_p, w_Ap, S, n, c.d_partials, c.d_ticket,
                       xrank(), d_status));
      count_launch();
    }
    launched++;
    if (launched > LAG) {
      const long long want = launched - LAG;
      const CgStatus* slot = h_status + (want % RING);
      long long spins = 0;
      while (slot->seq != want) {
        if ((++spins & 0xffff) == 0) {  // surface device faults instead of spinning forever
          cudaError_t e = cudaStreamQuery(c.stream);
          if (e != cudaSuccess && e != cudaErrorNotReady) return fail("cg_xx: %s", cudaGetErrorString(e));
          QB_CHECK(check_device_error());
        }
      }
      done = slot->done != 0;
    }
  }
  QB_CUDA(cudaGetLastError());
  double fin[S_N];
  QB_CHECK(fetch_scalars(S, fin, S_N));
  res->iterations = (int)fin[S_ITN];
  res->r2 = fin[S_R2];
  res->b2 = b2;
  if (keep) { keep->b2 = b2; keep->r2 = fin[S_R2]; keep->r2old = fin[S_R2OLD]; keep->iterations = (int)fin[S_ITN]; }
  return 0;
}

// =============================================================================================
// Multi-shift CG: CgmState.solve with cpNone (src/solvers/cgm.nim:84-316), new-solution branch,
// as driven by solveXX(s, xs, b, ms, ...) (src/physics/stagSolve.nim:296-345):
//   (A + sigma_k) xs[k] = b,  A = stagD2ee/oo(m0^2),  sigma_0 = 0, sigma_k = shifts[k].
// One iteration = 2 sweeps (Re<p,Ap> fused) + k_cgm_xr (x0/r update, |r|^2, then on the last
// thread the scalar recurrences zeta/alpha/beta of every shift, cgm.nim:252-266) + k_cgm_shift
// (ONE pass over r updating p0 and every xs[k], ps[k]).  All scalars live in device memory.
// =============================================================================================
constexpr int CGM_MAX = 64;
// per-shift device scalars: [0] zi, [1] zim1, [2] cx = alpha*zr, [3] cr = zip1, [4] cp = beta*zr^2, [5] sigma
struct CgmDev {
  double2* xs[CGM_MAX];
  double2* ps[CGM_MAX];
  double sc[CGM_MAX][6];
  double alphaim1, betaim1, beta;
  int continuing, nmass;
};
static CgmDev* d_cgm = nullptr;
static double2* w_ps = nullptr;
static size_t w_ps_elems = 0;

__global__ void k_cgm_init(double* S, CgStatus* st, CgmDev* M, double r2req, double maxits) {
  S[S_R2] = S[S_B2];                 // r = b, x = 0 (cgm.nim:166-172)
  S[S_R2OLD] = 1.0;
  S[S_R2STOP] = r2req * S[S_B2];
  S[S_ITN] = 0.0;
  S[S_MAXITS] = maxits;
  S[S_DONE] = (!(S[S_R2] > S[S_R2STOP]) || maxits <= 0.0) ? 1.0 : 0.0;   // :187 (maxits <= 0: our guard)
  S[S_SEQ] = 0.0;
  M->alphaim1 = -1.0; M->betaim1 = 0.0; M->beta = 0.0; M->continuing = 0;  // :207-208
  for (int m = 0; m < M->nmass; m++) { M->sc[m][0] = 1.0; M->sc[m][1] = 1.0; }
  st->itn = 0; st->done = (long long)S[S_DONE]; st->r2 = S[S_R2];
  __threadfence_system();
  st->seq = 0;
}

__global__ void __launch_bounds__(CB) k_cgm_xr(double2* __restrict__ x, double2* __restrict__ r,
                                                const double2* __restrict__ p, const double2* __restrict__ Ap,
                                                double* S, size_t n, double* partials, unsigned int* ticket,
                                                XRank* xr, CgStatus* st, CgmDev* M) {
  if (S[S_DONE] != 0.0) {
    if (blockIdx.x == 0 && threadIdx.x == 0) cg_publish(S, st);
    return;
  }
  const double qLAp = S[S_PAP];
  const double alpha = (qLAp != 0.0) ? S[S_R2] / qLAp : 0.0;     // cgm.nim:234-236
  double s = 0.0;
  for (size_t i = (size_t)blockIdx.x * CB + threadIdx.x; i < n; i += (size_t)gridDim.x * CB) {
    const double2 pv = p[i], av = Ap[i];
    double2 xv = x[i], rv = r[i];
    rv.x = fma(-alpha, av.x, rv.x); rv.y = fma(-alpha, av.y, rv.y);
    xv.x = fma(alpha, pv.x, xv.x); xv.y = fma(alpha, pv.y, xv.y);
    x[i] = xv; r[i] = rv;
    s = fma(rv.x, rv.x, s); s = fma(rv.y, rv.y, s);
  }
  if (grid_reduce<CB>(s, partials, ticket, S + S_R2NEW, xr)) {
    const double r2i = S[S_R2], r2ip1 = S[S_R2NEW];
    const double beta = (r2i != 0.0) ? r2ip1 / r2i : 0.0;        // :252-254
    S[S_ITN] += 1.0;
    const bool cont = (S[S_ITN] < S[S_MAXITS]) && (r2ip1 > S[S_R2STOP]);   // :256
    const double alphaim1 = M->alphaim1, betaim1 = M->betaim1;
    for (int m = 1; m < M->nmass; m++) {                         // :260-274
      double* c = M->sc[m];
      const double zi = c[0], zim1 = c[1], sg = c[5];
      double zip1d = alpha * betaim1 * (zim1 - zi);
      zip1d += zim1 * alphaim1 * (1.0 + sg * alpha);
      const double zip1 = (zip1d != 0.0) ? zi * zim1 * alphaim1 / zip1d : 0.0;
      const double zr = (zi != 0.0) ? zip1 / zi : 0.0;
      c[2] = alpha * zr; c[3] = zip1; c[4] = beta * zr * zr;
      if (cont) { c[1] = zi; c[0] = zip1; }
    }
    M->alphaim1 = alpha; M->betaim1 = beta; M->beta = beta; M->continuing = cont ? 1 : 0;   // :276
    S[S_R2OLD] = r2i;
    S[S_R2] = r2ip1;
    if (!cont) S[S_DONE] = 1.0;
    cg_publish(S, st);
  }
}

// After iteration `launch_no` really ran (S_ITN == launch_no): xs[k] += cx_k ps[k] for k >= 1 and, if the
// loop continues, p0 = r + beta p0, ps[k] = cr_k r + cp_k ps[k] (cgm.nim:258,271-273).  r is read once.
__global__ void __launch_bounds__(CB) k_cgm_shift(const double2* __restrict__ r, const double* S,
                                                   const CgmDev* __restrict__ M, size_t n, double launch_no) {
  if (S[S_ITN] != launch_no) return;
  const int nmass = M->nmass;
  const bool cont = M->continuing != 0;
  const double beta = M->beta;
  for (size_t i = (size_t)blockIdx.x * CB + threadIdx.x; i < n; i += (size_t)gridDim.x * CB) {
    const double2 rv = r[i];
    if (cont) {
      double2 pv = M->ps[0][i];
      pv.x = fma(beta, pv.x, rv.x); pv.y = fma(beta, pv.y, rv.y);
      M->ps[0][i] = pv;
    }
    for (int m = 1; m < nmass; m++) {
      const double cx = M->sc[m][2], cr = M->sc[m][3], cp = M->sc[m][4];
      double2 pv = M->ps[m][i], xv = M->xs[m][i];
      xv.x = fma(cx, pv.x, xv.x); xv.y = fma(cx, pv.y, xv.y);
      M->xs[m][i] = xv;
      if (cont) {
        pv.x = cr * rv.x + cp * pv.x; pv.y = cr * rv.y + cp * pv.y;
        M->ps[m][i] = pv;
      }
    }
  }
}

// ps[k] := r for every k (q := z ; ps[m] := p, cgm.nim:212-220)
__global__ void __launch_bounds__(CB) k_cgm_start(const double2* __restrict__ r, const CgmDev* __restrict__ M, size_t n) {
  for (size_t i = (size_t)blockIdx.x * CB + threadIdx.x; i < n; i += (size_t)gridDim.x * CB) {
    const double2 rv = r[i];
    for (int m = 0; m < M->nmass; m++) M->ps[m][i] = rv;
  }
}

static void cgm_free() {
  if (w_ps) cudaFree(w_ps);
  if (d_cgm) cudaFree(d_cgm);
  w_ps = nullptr; w_ps_elems = 0; d_cgm = nullptr;
}

int cgm_xx(const Gauge& g, int parity, double2* const* xs, const double2* b, double m2, const double* sigmas,
           int nmass, double r2req, int maxits, CgResult* res) {
  Context& c = ctx();
  if (nmass < 1 || nmass > CGM_MAX) return fail("multi-shift CG supports 1..%d masses, got %d", CGM_MAX, nmass);
  if (c.g.nranks > 1 && !c.p2p) return fail("multi-shift CG on several ranks needs the peer-memory transport (qexb200_p2p_attach)");
  QB_CHECK(cg_workspace());
  const size_t n = (size_t)3 * c.g.Vh;
  if (w_ps_elems < n * nmass) {
    if (w_ps) cudaFree(w_ps);
    w_ps = nullptr; w_ps_elems = 0;
    QB_CUDA(cudaMalloc(&w_ps, sizeof(double2) * n * nmass));
    w_ps_elems = n * nmass;
  }
  if (!d_cgm) QB_CUDA(cudaMalloc(&d_cgm, sizeof(CgmDev)));
  static CgmDev h;   // ~4 KB: keep it off the stack
  for (int m = 0; m < nmass; m++) {
    h.xs[m] = xs[m]; h.ps[m] = w_ps + (size_t)m * n;
    h.sc[m][5] = (m == 0) ? 0.0 : sigmas[m];                     // cgm.nim:107-110
  }
  h.nmass = nmass;
  double* S = c.d_scal;
  const int grid = cg_grid(n);
  QB_CUDA(cudaStreamSynchronize(c.stream));
  QB_CUDA(cudaMemcpyAsync(d_cgm, &h, sizeof h, cudaMemcpyHostToDevice, c.stream));
  for (int i = 0; i < RING; i++) h_status[i].seq = -1;

  QB_CHECK(blas_norm2(b, n, S + S_B2));
  QB_CHECK(allreduce_scalars(S + S_B2, 1));
  QB_CHECK(blas_copy(w_r, b, n));
  for (int m = 0; m < nmass; m++) QB_CHECK(blas_zero(xs[m], n));
  k_cgm_init<<<1, 1, 0, c.stream>>>(S, d_status, d_cgm, r2req, (double)maxits);
  k_cgm_start<<<grid, CB, 0, c.stream>>>(w_r, d_cgm, n);
  count_launch(2);
  double hs[2];
  QB_CHECK(fetch_scalars(S + S_B2, hs, 1));
  const double b2 = hs[0];
  res->b2 = b2;
  if (b2 == 0.0) { res->iterations = 0; res->r2 = 0.0; return 0; }

  double2* p0 = w_ps;
  const int LAG = 3;
  int launched = 0;
  bool done = h_status[0].done != 0;
  while (!done && launched < maxits) {
    QB_CHECK(apply_xx(g, parity, w_Ap, p0, m2, S + S_PAP, S + S_DONE));
    QB_CHECK(allreduce_scalars(S + S_PAP, 1));
    k_cgm_xr<<<grid, CB, 0, c.stream>>>(xs[0], w_r, p0, w_Ap, S, n, c.d_partials, c.d_ticket, xrank(), d_status, d_cgm);
    launched++;
    k_cgm_shift<<<grid, CB, 0, c.stream>>>(w_r, S, d_cgm, n, (double)launched);
    count_launch(2);
    if (launched > LAG) {
      const long long want = launched - LAG;
      const CgStatus* slot = h_status + (want % RING);
      long long spins = 0;
      while (slot->seq != want) {
        if ((++spins & 0xffff) == 0) {
          cudaError_t e = cudaStreamQuery(c.stream);
          if (e != cudaSuccess && e != cudaErrorNotReady) return fail("cgm_xx: %s", cudaGetErrorString(e));
          QB_CHECK(check_device_error());
        }
      }
      done = slot->done != 0;
    }
  }
  QB_CUDA(cudaGetLastError());
  double fin[S_N];
  QB_CHECK(fetch_scalars(S, fin, S_N));
  res->iterations = (int)fin[S_ITN];
  res->r2 = fin[S_R2];
  return 0;
}

}  // namespace qb
