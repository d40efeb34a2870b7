/* qex_b200.h -- C ABI of the B200-native staggered Dslash + CG backend for QEX.
 *
 * This is the drop-in boundary: exactly what a QEX `importc` wrapper binds, in
 * the same place and shape as QEX's existing QUDA backend.  Citations are
 * file:line in the reference tree (ctpeterson/qex).
 *
 *   plug point        SolverParams.backend / `case sp.backend` in solveXX
 *                     (src/solvers/solverBase.nim:6-14, src/physics/stagSolve.nim:57-132)
 *   ABI precedent     qudaInit/qudaInvert/qudaFinalize
 *                     (src/quda/quda_milc_interface.nim:75-146,289-295,
 *                      src/quda/qudaWrapperImpl.nim:71-261)
 *
 * Conventions
 *   - every entry point returns 0 on success, non-zero on error; the message is
 *     available from qexb200_last_error() (QEX glue turns it into qexError).
 *   - one host thread per rank calls the library, from outside `threads:`
 *     (like qudaSolveXX, src/quda/qudaWrapperImpl.nim:165).
 *   - one rank <-> one GPU.  Ranks are numbered as QEX numbers them:
 *     lexicographic in rankGeom with dimension 0 slowest (src/layout/qlayout.nim:94-113).
 *   - host field pointers are NATIVE QEX memory (`Field.s.data`,
 *     src/field/fieldET.nim:18-22,209): outer site e, then colour (row, col),
 *     then re/im, then the V SIMD lanes:
 *        colour vector  double[nSitesOuter][3][2][V]
 *        colour matrix  double[nSitesOuter][3][3][2][V]
 *     with QEX's site order (even outer sites first, src/layout/qlayout.nim:110-131).
 *     V and the lane geometry are given once by qexb200_set_host_layout().
 *   - parity / subset: 0 = even, 1 = odd, 2 = all.
 *   - the library never keeps a host pointer past the call.
 *   - there is no CPU fallback: without a CUDA device every compute entry point
 *     fails with an error.
 */
#ifndef QEX_B200_H
#define QEX_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define QEXB200_EVEN 0
#define QEXB200_ODD 1
#define QEXB200_ALL 2

typedef int qexb200_gauge_t; /* handle of a device-resident link set */
typedef int qexb200_field_t; /* handle of a device-resident colour-vector field */
typedef int qexb200_rng_t;   /* handle of a device-resident per-site random number field */

/* SolverParams, src/solvers/solverBase.nim:10-28 (inputs first, outputs after). */
typedef struct {
  double r2req;      /* stop when r2 <= r2req * |b|^2 (squared relative residual) */
  int maxits;
  int use_prev_soln; /* usePrevSoln: x on entry is the initial guess (qexb200_solve only) */
  int verbosity;
  int sloppy_solve;  /* sloppySolve: 0 SloppyNone, 1 SloppySingle, 2 SloppyHalf (solverBase.nim:8-9,15).  ACCEPTED BUT NOT
                        USED YET: this library has fp64 kernels only and solves in full fp64 whatever is asked -- what
                        QEX's own backend does too (only its QUDA backend maps the field to mixed precision,
                        src/quda/qudaWrapperImpl.nim:194-197).  sloppy_used (below) tells the caller. */
  /* outputs of THIS call (the Nim glue folds them into sp0 with addStats, solverBase.nim:75-83) */
  int calls;
  int iterations;
  int iterations_max;
  double seconds;
  double flops;      /* QEX flop model (g.len*4*72+60)*nEven*iterations, stagSolve.nim:92 */
  double r2;         /* solveXX: final recursive r2/b2; solve: true |b-Dx|^2/|b|^2 */
  int sloppy_used;   /* precision the inner solve actually ran in: always 0 (fp64) in this version */
} qexb200_solver_params;

/* ---- lifetime ---------------------------------------------------------- */

/* Replaces qudaInit + initCommsGridQuda (qudaWrapperImpl.nim:71-127).
 * latsize = physGeom, rankgeom = rankGeom (product = number of ranks).
 * nccl_unique_id: the 128-byte ncclUniqueId created on rank 0 with
 * qexb200_nccl_unique_id() and broadcast by the caller (Comm.broadcast,
 * src/comms/comms.nim:13); NULL when there is a single rank. */
int qexb200_init(const int latsize[4], const int rankgeom[4], int myrank, int device,
                 const void* nccl_unique_id);
int qexb200_nccl_unique_id(void* out128);
/* Peer-memory transport for the ranks of ONE box (NVLink / NVSwitch): every rank exports a 64-byte
 * handle of its receive window (cudaIpcMemHandle_t), the caller all-gathers the handles in rank
 * order (the same Comm the id above travels on, src/comms/comms.nim:13) and every rank attaches.
 * From then on the face-pack kernel stores halos straight into the neighbours' windows, the
 * boundary kernel waits on flags the neighbours raise, and the CG's scalar sums are completed
 * inside the reducing kernels: the startSB / QMP_wait / boundarySB2 sequence of
 * src/layout/shifts.nim:67-94,254-285 and the two QMP_sum_double per iteration (src/solvers/cg.nim:206,213)
 * without a host round trip.  Optional: without it the NCCL send/recv path is used; with it,
 * qexb200_init may be given a NULL nccl_unique_id. */
int qexb200_p2p_export(void* handle_out64);
int qexb200_p2p_attach(const void* handles_nranks_x_64);
int qexb200_p2p_active(void);
/* Replaces qudaFinalize (quda_milc_interface.nim:146). */
int qexb200_finalize(void);
const char* qexb200_last_error(void);
/* 1 if a CUDA device was found and the context is live. */
int qexb200_is_initialized(void);

/* Lane geometry of the caller's fields = Layout.innerGeom (src/layout/layoutX.nim:98-125).
 * {1,1,1,1} (the default) is a V=1 layout. */
int qexb200_set_host_layout(const int innergeom[4]);
/* Local (per-rank) number of sites, and of even sites. */
int qexb200_local_sites(int* nsites, int* neven);

/* ---- gauge links: Staggered.g (src/physics/stagD.nim:19-22,522-564) ------ */

/* nlinks = 4: naive one-hop, links[mu].  nlinks = 8: fat + Naik, links[2mu] one-hop,
 * links[2mu+1] three-hop, the order newStag3 builds (stagD.nim:552-564).
 * The caller has already applied setBC/stagPhase (stagD.nim:72-80,509-520) or uses
 * qexb200_gauge_rephase. */
int qexb200_gauge_create(int nlinks, qexb200_gauge_t* g);
int qexb200_gauge_load(qexb200_gauge_t g, const double* const* links);
int qexb200_gauge_destroy(qexb200_gauge_t g);
/* Resident handle for a caller that owns HOST link fields (the Nim glue of INTEGRATION.md section 2): the device copy
 * is cached against (links[0..nlinks) pointers, version).  Same pointers and same version: the handle made earlier is
 * returned and nothing is uploaded (*cached = 1).  Same pointers, new version: the links are uploaded again INTO the
 * same handle (*cached = 0).  New pointers: create + load; at most 8 sets are kept, the least recently used one is
 * destroyed.  `version` is the caller's update counter of these links -- QEX rewrites fermion links only through
 * smear -> setBC -> stagPhase (src/stagg_pv_hmc/staghmc_spv.nim:367-386, src/physics/stagD.nim:72-80), the glue bumps
 * its counter there.  What qudaSolveXX re-does on every call (src/quda/qudaWrapperImpl.nim:198-261) is paid once per
 * link update.  Handles obtained here are owned by the cache: do not pass them to qexb200_gauge_destroy.
 * Guard against a forgotten bump: on every call the library also fingerprints a strided sample of the host links
 * (1024 doubles per field, a few KB) and uploads again when it differs from the one taken at upload time, even if
 * the version is unchanged -- this catches whole-field rewrites (smear, rephase, a new configuration, a freed set
 * whose addresses were reused), not a change confined to sites between the samples, so the counter remains the
 * contract.  When link arrays are freed, call qexb200_gauge_invalidate with their pointers (or use a version that
 * is never repeated across distinct link objects). */
int qexb200_gauge_cached(const double* const* links, int nlinks, unsigned long long version,
                         qexb200_gauge_t* g, int* cached);
/* Drops the cached device copy keyed on these host pointers, if any (no error when there is none).  For the glue's
 * finaliser of a Staggered object: qudaSolveXX needs nothing like it because it uploads on every call
 * (src/quda/qudaWrapperImpl.nim:198-261). */
int qexb200_gauge_invalidate(const double* const* links, int nlinks);
/* Link storage the sweeps read.  want = 18: all 18 reals of every link (default).  want = 12: each
 * link class (one-hop, three-hop) whose matrices are s*SU(3) with a single |s| -- naive links with
 * staggered phases and BC (+-SU(3)), Naik links -- is additionally held as rows 0,1 and the third
 * row is rebuilt in registers (96 instead of 144 bytes per link, the bandwidth lever of this
 * HBM-bound stencil); classes that do not qualify (fat7-smeared links) stay at 18.  The library
 * verifies the form on the device (rebuilt matrix within 2e-14 |s|) and reports per class what it
 * uses in got[2] (12 or 18) and the measured deviation in dev[2].  want = 1812 / 1218 restricts
 * the request to the three-hop / one-hop class (1812: fat links 18, Naik links 12 -- what a real
 * HISQ or asqtad link set allows).  want = 14 (or 1814, 1418, 1214 ...): rows 0,1 plus ONE complex
 * factor phi per link, third row = phi conj(row0 x row1) (112 bytes per link) -- for links that are
 * c*U(3) with a determinant phase that changes from site to site: projectU'd smeared links (nHYP,
 * src/gauge/hypsmear.nim:49-162) and the Naik products of the HISQ W links (src/gauge/fat7l.nim:156-167).
 * Same on-device verification.  The request is sticky: it is re-applied after every
 * qexb200_gauge_load / qexb200_gauge_rephase and after every device smear that writes the set. */
int qexb200_gauge_recon(qexb200_gauge_t g, int want, int got[2], double dev[2]);
/* Staggered.rephase on the device copy: setBC then stagPhase with phases {8,9,11,0}
 * (stagD.nim:72-80,509-522; gauge/gaugeUtils.nim:124-131).  One-hop links only. */
int qexb200_gauge_rephase(qexb200_gauge_t g);

/* ---- device link pipeline ------------------------------------
 * The smearing steps QEX runs on the CPU before every solve of an HMC force step
 * (src/stagg_pv_hmc/staghmc_spv.nim:367-386; src/physics/hisqLinks.nim:32-39), on the resident links:
 * thin links are uploaded once with qexb200_gauge_load, smeared, rephased and solved on without
 * leaving the GPU.  All handles here hold 4 link fields unless stated.  On a lattice split over ranks the
 * peer-memory transport is required (qexb200_p2p_attach BEFORE the gauges are created): link-shaped fields then
 * live in an arena of the peer window (QEXB200_ARENA_FIELDS matrix fields, default 64) and off-rank neighbours
 * are read over NVLink; with NCCL only these return an error (no silent host path). */
typedef struct {   /* Fat7lCoefs, src/gauge/fat7l.nim:5-10 */
  double one_link, three_staple, five_staple, seven_staple, lepage;
} qexb200_fat7l_coefs;
typedef struct {   /* HisqCoefs, src/physics/hisqLinks.nim:4-8 */
  qexb200_fat7l_coefs fat7first, fat7second;
  double naik;
} qexb200_hisq_coefs;
typedef struct {   /* HypCoefs, src/gauge/hypsmear.nim:16-19 */
  double alpha1, alpha2, alpha3;
} qexb200_hyp_coefs;
/* HisqCoefs.init (hisqLinks.nim:17-24) */
void qexb200_hisq_coefs_init(qexb200_hisq_coefs* c);
/* makeImpLinks(fl, gf, coef, ll, gfLong, naik) (fat7l.nim:78-162).  naik == 0: ll / gfLong are
 * ignored (pass -1), the 4-argument overload of fat7l.nim:164-165. */
int qexb200_gauge_makeImpLinks(qexb200_gauge_t fl, qexb200_gauge_t gf, const qexb200_fat7l_coefs* coef,
                               qexb200_gauge_t ll, qexb200_gauge_t gfLong, double naik);
/* HisqCoefs.smear(g, fl, ll) (hisqLinks.nim:32-43): fat7 -> projectU -> fat7 + Lepage + Naik */
int qexb200_gauge_smear_hisq(qexb200_gauge_t fl, qexb200_gauge_t ll, qexb200_gauge_t g,
                             const qexb200_hisq_coefs* coef);
/* HypCoefs.smear(g, fl) (hypsmear.nim:49-162,269-273): three levels of projected staples */
int qexb200_gauge_smear_hyp(qexb200_gauge_t fl, qexb200_gauge_t g, const qexb200_hyp_coefs* coef);
/* smearGetForce (hypsmear.nim:49-162): like qexb200_gauge_smear_hyp, and keeps the projected and unprojected
 * fields of the three levels in a state handle (52 matrix fields).  The state refers to g: g must stay unchanged
 * and alive until the state is destroyed (hypsmear.nim:51-53). */
typedef int qexb200_hyp_t;
int qexb200_hyp_smear_get_force(qexb200_gauge_t fl, qexb200_gauge_t g, const qexb200_hyp_coefs* coef,
                                qexb200_hyp_t* state);
/* smearedForce(f, chain) (hypsmear.nim:163-262): the chain rule through projectU (projectUderiv,
 * src/maths/matrixFunctions.nim:321-357) and the staples (symStapleDeriv, src/gauge/smearutil.nim:24-54) of the three
 * levels: chain = derivative with respect to the smeared links, f = with respect to the thin links.  f may be the
 * same handle as chain (f.smeared_force(f), src/stagg_pv_hmc/staghmc_spv.nim:745).  On a split lattice it needs
 * QEXB200_ARENA_FIELDS >= 100. */
int qexb200_hyp_smeared_force(qexb200_hyp_t state, qexb200_gauge_t f, qexb200_gauge_t chain);
int qexb200_hyp_destroy(qexb200_hyp_t state);
/* fat7lDeriv(deriv, gauge, mid, coef, llderiv, llgauge, llmid, naik) (src/gauge/fat7lderiv.nim:122-335): the derivative
 * of makeImpLinks, mid / llmid = chain with respect to the fat / long links; deriv and llderiv ACCUMULATE (zero them
 * first where the reference passes fresh fields).  naik == 0: the ll* handles are ignored (pass -1). */
int qexb200_gauge_fat7lDeriv(qexb200_gauge_t deriv, qexb200_gauge_t gauge, qexb200_gauge_t mid,
                             const qexb200_fat7l_coefs* coef, qexb200_gauge_t llderiv, qexb200_gauge_t llgauge,
                             qexb200_gauge_t llmid, double naik);
/* HisqCoefs.smearGetForce(u, su, sul) (src/gauge/hisqsmear.nim:51-84): smears like qexb200_gauge_smear_hisq and keeps
 * the first-level fat links and their projection; smearedForce(dsdu, dsdsu, dsdsul) then runs second fat7 + Naik
 * derivative -> projectUderiv -> first fat7 derivative, ACCUMULATING into dsdu (examples/hisqhmc.nim:496-530). */
typedef int qexb200_hisq_t;
int qexb200_hisq_smear_get_force(qexb200_gauge_t su, qexb200_gauge_t sul, qexb200_gauge_t u,
                                 const qexb200_hisq_coefs* coef, qexb200_hisq_t* state);
int qexb200_hisq_smeared_force(qexb200_hisq_t state, qexb200_gauge_t dsdu, qexb200_gauge_t dsdsu, qexb200_gauge_t dsdsul);
int qexb200_hisq_destroy(qexb200_hisq_t state);
/* projTAH(f, g, adj) (src/stagg_pv_hmc/staghmc_spv_gforce.nim:256-291): f[mu] := projectTAH(f[mu] g[mu]^+)
 * (adj = 0, matter force) or projectTAH(g[mu] f[mu]^+) (adj != 0, gauge force); projectTAH = traceless
 * anti-Hermitian part (src/maths/matrixFunctions.nim:375-380). */
int qexb200_gauge_projTAH(qexb200_gauge_t f, qexb200_gauge_t g, int adj);
/* dst[mu][x] = projectU(src[mu][x]) (src/maths/matrixFunctions.nim:301-312); same link count */
int qexb200_gauge_projectU(qexb200_gauge_t dst, qexb200_gauge_t src);
/* plaq (src/gauge/gaugeUtils.nim:213-284) of the one-hop links: pl[mu(mu-1)/2 + nu], nu < mu */
int qexb200_gauge_plaq(qexb200_gauge_t g, double pl[6]);
/* newStag3(fl, ll) (stagD.nim:552-564): out (8 link fields) = {fl[0], ll[0], fl[1], ll[1], ...} */
int qexb200_gauge_join3(qexb200_gauge_t out, qexb200_gauge_t fl, qexb200_gauge_t ll);
/* the inverse: fl[mu] = in[2 mu], ll[mu] = in[2 mu + 1] -- e.g. the one-hop and three-hop outer products of
 * examples/hisqhmc.nim:509-514 (f1, f3), formed in one pass by qexb200_dev_outer on an 8-field set */
int qexb200_gauge_split3(qexb200_gauge_t fl, qexb200_gauge_t ll, qexb200_gauge_t in);
/* device links -> native QEX host memory, links[k] as in qexb200_gauge_load */
int qexb200_gauge_download(qexb200_gauge_t g, double* const* links);

/* ---- operator on HOST fields (each call: H2D, kernels, D2H) --------------- */

/* stagD2 (stagD.nim:348-395): r[p] = a*r + b*x + sum_k [U_k x(+h_k) - U_k^+ x(-h_k)] */
int qexb200_stagD2(qexb200_gauge_t g, int parity, double* r, const double* x, double a, double b);
/* stagD (stagD.nim:405-409): r[p] = a*r + m*x + sc*D*x */
int qexb200_stagD(qexb200_gauge_t g, int parity, double* r, const double* x, double m, double sc,
                  double a);
/* D / Ddag (stagD.nim:566-571), both parities */
int qexb200_D(qexb200_gauge_t g, double* r, const double* x, double m);
int qexb200_Ddag(qexb200_gauge_t g, double* r, const double* x, double m);
/* peqDdag (stagD.nim:572-574): r += Ddag x;  eoReduce (:575-580): r.even = (Ddag b).even */
int qexb200_peqDdag(qexb200_gauge_t g, double* r, const double* x, double m);
int qexb200_eoReduce(qexb200_gauge_t g, double* r, const double* b, double m);
/* stagDb (stagD.nim:425-427): stagD2 with b = m/(0.5 sc), i.e. (m x + sc D x)/(0.5 sc) without the rescaling pass */
int qexb200_stagDb(qexb200_gauge_t g, int parity, double* r, const double* x, double m, double sc);
/* stagD1 (minus = 0) / stagD1x (minus != 0) (stagD.nim:414-422): r[p] = +- sum_k [U_k x(+h_k) - U_k^+ x(-h_k)] */
int qexb200_stagD1(qexb200_gauge_t g, int parity, double* r, const double* x, int minus);
/* stagD2ee / stagD2oo (stagD.nim:433-469): r[p] = 4 m2 x - (2D)(2D) x */
int qexb200_stagD2xx(qexb200_gauge_t g, int parity, double* r, const double* x, double m2);
/* eoReconstruct (stagD.nim:582-586): r.odd = (b.odd - D_oe r.even)/m */
int qexb200_eoReconstruct(qexb200_gauge_t g, double* r, const double* b, double m);

/* ---- solvers on HOST fields ---------------------------------------------- */

/* solveXX / solveEE / solveOO with the device backend: same contract as qudaSolveEE/OO
 * (qudaWrapperImpl.nim:165-267, stagSolve.nim:57-138): r := 0, CG on
 * stagD2ee/oo(m*m) r = x restricted to `parity`. */
int qexb200_solveXX(qexb200_gauge_t g, int parity, double* r, const double* x, double m,
                    qexb200_solver_params* sp);
/* solve (stagSolve.nim:224-294): full-lattice D x = b with even-odd preconditioning,
 * reconstruction and the true-residual outer loop. */
int qexb200_solve(qexb200_gauge_t g, double* x, const double* b, double m,
                  qexb200_solver_params* sp);

/* ---- multi-shift solves (src/solvers/cgm.nim:84-316, src/physics/stagSolve.nim:296-446) ---- */

/* solveXX(s, xs, b, ms, sp0, subset): (A + sigma_k) xs[k] = b on `parity`, A = stagD2ee/oo(shifts[0]^2),
 * sigma_0 = 0, sigma_k = shifts[k] for k > 0 (the caller passes shifts[0] = m_0, shifts[k] =
 * 4 (m_k^2 - m_0^2), stagSolve.nim:391-394).  xs[k] is zeroed on the subset, the other parity is
 * left alone.  sp->iterations = iterations of the unshifted system; sp->r2 = 0 like the reference. */
int qexb200_solveXX_multi(qexb200_gauge_t g, int parity, double* const* xs, const double* b,
                          const double* shifts, int nmass, qexb200_solver_params* sp);
/* solve(s, xs, b, ms, sp0): xs[k] = Ddag(m_k) (D Ddag)(m_k)^-1 b through the even/odd multi-shift
 * solve with the true-residual loop on mass ms[0] (stagSolve.nim:347-446). */
int qexb200_solve_multi(qexb200_gauge_t g, double* const* xs, const double* b, const double* ms,
                        int nmass, qexb200_solver_params* sp);

/* ---- device-resident fields (callers that keep vectors on the GPU) ------- */

int qexb200_field_create(qexb200_field_t* f);
int qexb200_field_destroy(qexb200_field_t f);
int qexb200_field_upload(qexb200_field_t f, const double* host);
int qexb200_field_download(qexb200_field_t f, double* host);
int qexb200_field_zero(qexb200_field_t f);
/* norm2 / redot on a subset (src/field/fieldET.nim:605-625,704-724), summed over ranks */
int qexb200_field_norm2(qexb200_field_t f, int subset, double* out);
int qexb200_field_redot(qexb200_field_t f, qexb200_field_t g, int subset, double* out);
/* y = a x + b y on a subset: the assignments `y := x`, `y += a*x`, `y := a*x + b*y` of the field expression
 * templates (src/field/fieldET.nim:547-599) for callers that keep their vectors on the GPU. */
int qexb200_field_axpby(qexb200_field_t y, double a, qexb200_field_t x, double b, int subset);

/* |x|^2 of a device array the CALLER owns, computed without allocating anything: no cudaMalloc / cudaFree, no
 * library context (qexb200_init is not needed), no torch.  It has no counterpart in the reference; it is the norm2
 * kernel above (qexb200_field_norm2) with its reduction scratch -- per-block partial sums and a completion counter --
 * moved from the context's cudaMalloc'd buffers into a static device array inside the library: 128.25 KB that every
 * GPU on which the library is loaded holds until the process exits.
 *
 * It is meant for callers that cannot allocate device memory when they need the number: runs sized to fill the GPU
 * (resident link sets of the largest lattices, or a device whose free memory belongs to another backend's pool or
 * another process), where a late cudaMalloc fails mid-trajectory; sections in which the allocator's implicit
 * device-wide synchronisation would stall kernels in flight on other streams; and monitoring code that only holds
 * raw device pointers from its own allocator and never creates a qexb200 context.
 *
 *   d_x     device pointer, 16-byte aligned, nreals doubles (a colour-vector field or any part of one)
 *   nreals  0 .. QEXB200_NOALLOC_MAX_REALS (1 GiB of doubles: one full-lattice vector at 64^4, one parity at
 *           64^3 x 128).  More is refused with an error -- split the array and add the pieces.
 *   d_out   device pointer to ONE double, written by the kernel
 *   stream  the caller's cudaStream_t (NULL = the default stream).  The call only enqueues (a 256-byte memset of the
 *           scratch's counter, then the kernel); *d_out is valid once the stream has reached that point.
 *
 * The sum is formed in a fixed order that depends on nreals alone (fixed 8192-real tiles, partials added in index
 * order), so it is reproducible bit for bit from call to call and from GPU to GPU; it differs from
 * qexb200_field_norm2 only by the rounding of a different summation order.
 *
 * Set-up: the FIRST call on each device resolves the device address of the static array (cudaGetSymbolAddress, with
 * that device current) and remembers it; make that first call on every device before starting any stream capture.
 * The device that is current at the call must be the one d_x, d_out and stream belong to.
 * Sharing: all calls on one GPU use the same scratch.  Calls ordered on one stream are safe; two calls that can
 * overlap on the GPU (different streams, or different host threads, of one process without ordering between them)
 * race on it and return garbage -- serialise them. */
#define QEXB200_NOALLOC_MAX_REALS 134217728LL
int qexb200_norm2_noalloc(const double* d_x, long long nreals, double* d_out, void* stream);

int qexb200_dev_stagD2(qexb200_gauge_t g, int parity, qexb200_field_t r, qexb200_field_t x,
                       double a, double b);
int qexb200_dev_stagD(qexb200_gauge_t g, int parity, qexb200_field_t r, qexb200_field_t x,
                      double m, double sc, double a);
int qexb200_dev_D(qexb200_gauge_t g, qexb200_field_t r, qexb200_field_t x, double m);
int qexb200_dev_Ddag(qexb200_gauge_t g, qexb200_field_t r, qexb200_field_t x, double m);
int qexb200_dev_stagD2xx(qexb200_gauge_t g, int parity, qexb200_field_t r, qexb200_field_t x,
                         double m2);
int qexb200_dev_peqDdag(qexb200_gauge_t g, qexb200_field_t r, qexb200_field_t x, double m);
int qexb200_dev_eoReduce(qexb200_gauge_t g, qexb200_field_t r, qexb200_field_t b, double m);
int qexb200_dev_solveXX(qexb200_gauge_t g, int parity, qexb200_field_t r, qexb200_field_t x,
                        double m, qexb200_solver_params* sp);
int qexb200_dev_solve(qexb200_gauge_t g, qexb200_field_t x, qexb200_field_t b, double m,
                      qexb200_solver_params* sp);
int qexb200_dev_solveXX_multi(qexb200_gauge_t g, int parity, const qexb200_field_t* xs, qexb200_field_t b,
                              const double* shifts, int nmass, qexb200_solver_params* sp);
int qexb200_dev_solve_multi(qexb200_gauge_t g, const qexb200_field_t* xs, qexb200_field_t b,
                            const double* ms, int nmass, qexb200_solver_params* sp);

/* ---- gauge import straight into the device layout --------------------------------
 * loadGauge (src/gauge/gaugeUtils.nim:87-106) through Reader.read (src/io/readerQiolite.nim:108-165): the
 * rank's hypercube of the record goes to the GPU as raw bytes; byte swap, float -> double and the
 * permutation into the device layout are one kernel.  4 link fields. */
/* getFileLattice (readerQiolite.nim:11-17); pure host, needs no GPU */
int qexb200_scidac_lattice(const char* filename, int dims[4]);
/* First lattice record of a SciDAC (or ILDG) file -> g.  sums (may be NULL) = {file suma, file sumb, this rank's
 * partial suma, partial sumb}: a single rank verifies the checksum itself (mismatch = error); several ranks
 * each read their own hypercube and the caller XORs the partials over the ranks. */
int qexb200_gauge_read_scidac(qexb200_gauge_t g, const char* filename, unsigned int sums[4]);
/* Site-major host array [local site, lexicographic x0 fastest][mu][3][3][re,im] of 4- or 8-byte reals, e.g. the
 * MILC-order arrays the QUDA wrapper builds (src/quda/qudaWrapperImpl.nim:216-240) or a reader's raw record. */
int qexb200_gauge_load_sitemajor(qexb200_gauge_t g, const void* data, int precision_bytes, int big_endian);

/* ---- fermion-force outer products (src/physics/stagD.nim:588-664) -----------------
 * Link-shaped outputs live in gauge handles with the operator's link count (4, or 8 with the three-hop
 * fields at 2 mu + 1).  Split lattices: as the link pipeline above (peer-memory transport). */
int qexb200_gauge_zero(qexb200_gauge_t g);            /* f[mu] := 0 */
int qexb200_gauge_scale(qexb200_gauge_t g, double c); /* g[mu] *= c (src/hmc/agradOps.nim:513-520) */
/* stagDeriv(s, f, x) (stagD.nim:635-664): f[mu](e) += x(e) (x) x(e+mu)^+, f[mu](o) -= x(o) (x) x(o+mu)^+,
 * then s.rephase f.  4 link fields. */
int qexb200_dev_stagDeriv(qexb200_gauge_t s, qexb200_gauge_t f, qexb200_field_t x);
int qexb200_stagDeriv(qexb200_gauge_t s, double* const* f, const double* x);
/* stagD2deriv(s, g, c, x) (stagD.nim:588-633) = (d/dg) redot[(2D) x, c]:
 * g[k] += c (x) x(+h_k)^+ - x (x) c(+h_k)^+ on all sites. */
int qexb200_dev_stagD2deriv(qexb200_gauge_t s, qexb200_gauge_t g, qexb200_field_t c, qexb200_field_t x);
int qexb200_stagD2deriv(qexb200_gauge_t s, double* const* g, const double* c, const double* x);
/* The force loop of src/stagg_pv_hmc/staghmc_spv.nim:826-850:
 * f[k](x) (+)= (x even ? scale_even : scale_odd) * a(x) (x) b(x+h_k)^+; accumulate == 0 overwrites. */
int qexb200_dev_outer(qexb200_gauge_t f, qexb200_field_t a, qexb200_field_t b, double scale_even,
                      double scale_odd, int accumulate);

/* ---- CgState as an object and solveEO (src/solvers/cg.nim:13-53,256-283; src/physics/stagSolve.nim:17-38) ----- */

typedef int qexb200_cg_t;    /* handle of a restartable CG state (CgState[T]: r, p and the recurrence scalars) */
/* solveEO(s, r, x, m, sp): cgSolve(r, x, stagD2ee(m^2)) on the even subset.  Unlike solveXX, r is NOT zeroed: its
 * even part on entry is the initial guess (r0 = x - A r is formed by the first call of CgState.solve, cg.nim:146-149).
 * sp.r2 = final recursive r2/b2. */
int qexb200_solveEO(qexb200_gauge_t g, double* r, const double* x, double m, qexb200_solver_params* sp);
int qexb200_dev_solveEO(qexb200_gauge_t g, qexb200_field_t r, qexb200_field_t x, double m, qexb200_solver_params* sp);
/* newCgState(x, b) bound to A = stagD2ee / stagD2oo of gauge g (parity QEXB200_EVEN / QEXB200_ODD).  The fields are
 * referenced, not copied; x holds the initial guess. */
int qexb200_cg_create(qexb200_gauge_t g, int parity, qexb200_field_t x, qexb200_field_t b, qexb200_cg_t* cg);
/* CgState.reset (cg.nim:21-27): the next solve starts over from the current x. */
int qexb200_cg_reset(qexb200_cg_t cg);
/* CgState.solve(op, sp) (cg.nim:55-254).  First call after create/reset: b2, r = b - A x.  Later calls CONTINUE the
 * same Krylov sequence (r, p, r2old, iteration counter carry over): raising sp.maxits or tightening sp.r2req and
 * calling again gives exactly the iterates of one uninterrupted solve.  sp.maxits and sp.iterations are cumulative,
 * as in the reference (while itn < maxits; sp.finalIterations = state.iterations). */
int qexb200_cg_solve(qexb200_cg_t cg, double mass, qexb200_solver_params* sp);
int qexb200_cg_destroy(qexb200_cg_t cg);

/* ---- per-site random number field: the input generator of the reference's tests and benchmarks ---------- */

/* newRNGField(RngMilc6, lo, seed) (src/rng/distributionUtils.nim:306-331): one RngMilc6 (src/rng/milcrng.nim:92-154)
 * per site, seeded with the GLOBAL lexicographic site index, so the fields do not depend on the rank geometry.  The
 * integer streams are bit-identical with QEX's; a deviate (milcrng.nim:158-188) is float32, as in QEX, but is formed
 * with the device's log/cos: about one deviate in 1e8 differs from QEX's by one float32 ulp (the double result sits on
 * a float32 rounding boundary and the two libms disagree in its last bit).  The 8^4 plaquettes of
 * tests/reprod/trandgauge.nim:17 are reproduced to sum diff^2 = 9e-32 (reference bar 1e-30). */
int qexb200_rng_create(unsigned long long seed, qexb200_rng_t* r);
int qexb200_rng_destroy(qexb200_rng_t r);
/* g.random r (src/gauge/gaugeUtils.nim:1348-1355,1424-1429): field by field, 18 deviates per site, projectSU. */
int qexb200_gauge_random(qexb200_gauge_t g, qexb200_rng_t r);
/* p.randomTAH r (gaugeUtils.nim:1353-1383,1452-1454): 8 deviates per site and field. */
int qexb200_gauge_randomTAH(qexb200_gauge_t g, qexb200_rng_t r);
/* g.unit (gaugeUtils.nim:1448-1450). */
int qexb200_gauge_unit(qexb200_gauge_t g);
/* reunit(g) of the HMC drivers (src/examples/staghmc_sh.nim:247-257): checkSU, projectSU on every link
 * (src/maths/matrixFunctions.nim:359-370), checkSU again.  dev[4] = {avg, max} deviation before and {avg, max} after,
 * normalised like checkSU (src/gauge/gaugeUtils.nim:1407-1422).  g.random leaves links unitary to ~1e-11 (closed-form
 * eigs3 on a gaussian matrix); after reunit they are unitary to ~1e-15 and qualify for the 12-real storage. */
int qexb200_gauge_reunit(qexb200_gauge_t g, double dev[4]);
/* v.gaussian r (src/rng/distributionUtils.nim:65-97): 6 deviates per site. */
int qexb200_field_gaussian(qexb200_field_t f, qexb200_rng_t r);

/* ---- halo plan (pure host arithmetic, usable without a GPU) ---------------- */

/* One entry per direction split over ranks: what this rank sends each parity sweep.  It is the
 * counterpart of the per-shift message tables QEX builds in makeShiftMultiSubQ / prepareShiftBufsQ
 * (src/layout/qshifts.nim:23-94,355-533).  To rank_minus go `depth` face layers of the source
 * vector (n_f complex numbers at off_f); to rank_plus go the U^+ x products of the last layers
 * (n_b at off_b), formed at the source like stagD.nim:359-361.  Receives mirror the sends. */
typedef struct {
  int mu, depth, faceh, rank_plus, rank_minus;
  long long off_f, n_f, off_b, n_b; /* in complex doubles */
} qexb200_halo_dir;
int qexb200_plan_halo(const int latsize[4], const int rankgeom[4], int myrank, int nlinks,
                      qexb200_halo_dir out[4], int* ndirs, long long* total_elems);

/* Peer table of the multi-rank link pipeline (pure host arithmetic): peer_of[nb] = rank owning the sites that are
 * disp_d in {-1, 0, +1} rank-grid steps away in direction d, nb = sum_d (disp_d + 1) 3^d (40 = this rank), periodic,
 * ranks numbered as in qexb200_init.  A site whose neighbour lies there reads it from that rank's arena. */
int qexb200_plan_peers(const int latsize[4], const int rankgeom[4], int myrank, unsigned char peer_of[81]);
/* Host arithmetic of the multi-rank sweep's block -> tile table (no CUDA): tiles of 128 consecutive checkerboard
 * sites of parity `parity`, those without an off-rank hop first, the face tiles last; entry = tile | need << 24,
 * need bit 2 mu: the tile reads the + neighbour's x layers in mu, bit 2 mu + 1: the - neighbour's U^+ x products
 * (the remote part of startSB / boundarySB2, src/layout/shifts.nim:67-94,254-285).  out holds ceil(Vh/128) words. */
int qexb200_plan_tiles(const int latsize[4], const int rankgeom[4], int myrank, int nlinks, int parity,
                       unsigned int* out, int* ntiles);
/* Diagnostic of the peer-memory transport: on = 1 arms a ring of per-launch role timestamps (%globaltimer, ns) inside
 * the multi-rank sweep -- 1024 launches x 8 words {first block start, last pack block end, last interior tile end,
 * first face tile past its flag wait, last face tile end, longest flag wait, sum of flag waits, face tiles that
 * waited} -- followed by 8 words for the cross-rank sums {ns inside, calls, longest}.  on = 0 copies up to nwords
 * words out and disarms.  The counterpart of the reference's per-phase timers tic/toc in stagD.nim:355-393. */
int qexb200_debug_role_times(int on, unsigned long long* out, int nwords);

/* ---- measurement helpers (CUDA events on the library's compute stream) ---- */

int qexb200_sync(void);
int qexb200_timer_start(void);
int qexb200_timer_stop(float* milliseconds); /* synchronises */
/* number of kernel launches issued by the library since the last call with reset != 0 */
long long qexb200_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* QEX_B200_H */
