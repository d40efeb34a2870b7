// Host <-> device field re-layout.
//
// The caller hands over NATIVE QEX memory (see include/qex_b200.h): outer site
// e, colour indices, re/im, then V SIMD lanes; QEX's site order is defined by
// layoutIndexQ (src/layout/qlayout.nim:110-131).  The raw bytes are copied to the
// device as they are (one contiguous PCIe transfer) and re-laid-out there by a
// gather kernel into the checkerboarded SoA layout of common.cuh, using a site
// permutation table built once per (geometry, innerGeom).  This replaces the
// per-call CPU loops of the QUDA wrapper (src/quda/qudaWrapperImpl.nim:198-261).
#include "stencil.cuh"

namespace qb {

// QEX single-site index i = outer*V + lane of the site with GLOBAL coordinates xg,
// for a layout with local extents X, lane geometry `inner` (qlayout.nim:110-131).
// Also returns the owning rank (lexicographic, dimension 0 slowest, qlayout.nim:94-113).
static void qex_site_index(const int* xg, const Geom& g, const int* inner, int* rank, int* index) {
  int outer[4], nOuter = 1, V = 1;
  for (int d = 0; d < 4; d++) { outer[d] = g.X[d] / inner[d]; nOuter *= outer[d]; V *= inner[d]; }
  // which direction carries the checkerboard twist of the lanes (layoutSetupQ, qlayout.nim:10-64)
  int nOddSplit = 0, cbDir = -1;
  for (int d = 0; d < 4; d++) {
    if (inner[d] > 1 && (outer[d] & 1)) nOddSplit++;
    if (inner[d] == 1 && !(outer[d] & 1)) cbDir = d;
  }
  int cbShift = 0;
  if (nOddSplit == 0) cbDir = 0;
  else cbShift = outer[cbDir] / 2;

  int r = 0, lane = 0, laneStride = 1, twist = 0, par = 0;
  int xo[4];
  for (int d = 0; d < 4; d++) {
    const int rcd = xg[d] / g.X[d];
    const int xl = xg[d] - rcd * g.X[d];
    const int ld = xl / outer[d];          // lane coordinate
    xo[d] = xl - ld * outer[d];            // outer coordinate
    r = r * g.G[d] + rcd;
    lane += ld * laneStride; laneStride *= inner[d];
    twist += ld * outer[d];
    par += xg[d];
  }
  if (twist & 1) xo[cbDir] = (xo[cbDir] + cbShift) % outer[cbDir];
  const int olex = ((xo[3] * outer[2] + xo[2]) * outer[1] + xo[1]) * outer[0] + xo[0];
  const int e = (par & 1) ? (olex + nOuter) / 2 : olex / 2;
  *rank = r;
  *index = e * V + lane;
}

int ensure_perm() {
  Context& c = ctx();
  int V = c.inner[0] * c.inner[1] * c.inner[2] * c.inner[3];
  if (c.d_perm && c.permV == V) return 0;
  const Geom& g = c.g;
  for (int d = 0; d < 4; d++)
    if (g.X[d] % c.inner[d] != 0) return fail("innerGeom[%d]=%d does not divide local extent %d", d, c.inner[d], g.X[d]);
  {
    int nOddSplit = 0, cbDir = -1;
    for (int d = 0; d < 4; d++) {
      const int o = g.X[d] / c.inner[d];
      if (c.inner[d] > 1 && (o & 1)) nOddSplit++;
      if (c.inner[d] == 1 && !(o & 1)) cbDir = d;
    }
    if (nOddSplit && cbDir < 0) return fail("QEX cannot lay out this innerGeom (no direction for the lane checkerboard)");
  }
  std::vector<int> perm((size_t)g.V);
  int bad = 0;
  for (int ds = 0; ds < g.V; ds++) {
    const int par = ds / g.Vh, idx = ds - par * g.Vh;
    int t = idx, x[4], xg[4];
    const int xh = g.X[0] / 2;
    int x0h = t % xh; t /= xh;
    x[1] = t % g.X[1]; t /= g.X[1];
    x[2] = t % g.X[2]; x[3] = t / g.X[2];
    x[0] = 2 * x0h + ((par + x[1] + x[2] + x[3]) & 1);
    for (int d = 0; d < 4; d++) xg[d] = g.rc[d] * g.X[d] + x[d];
    int r, i;
    qex_site_index(xg, g, c.inner, &r, &i);
    if (r != g.rank || i < 0 || i >= g.V) bad++;
    perm[ds] = i;
  }
  if (bad) return fail("host layout map is inconsistent (%d sites map off-rank)", bad);
  if (!c.d_perm) QB_CUDA(cudaMalloc(&c.d_perm, sizeof(int) * (size_t)g.V));
  QB_CUDA(cudaMemcpy(c.d_perm, perm.data(), sizeof(int) * (size_t)g.V, cudaMemcpyHostToDevice));
  c.permV = V;
  return 0;
}

static int ensure_stage(size_t bytes) {
  Context& c = ctx();
  if (c.stage_bytes >= bytes) return 0;
  if (c.d_stage) QB_CUDA(cudaFree(c.d_stage));
  c.d_stage = nullptr; c.stage_bytes = 0;
  QB_CUDA(cudaMalloc(&c.d_stage, bytes));
  c.stage_bytes = bytes;
  return 0;
}

constexpr int LB = 256;

// NC complex components per site (3: vector, 9: matrix).  dst[comp][ds] <- raw QEX memory.
template <int NC>
__global__ void __launch_bounds__(LB) k_gather(double2* __restrict__ dst, const double* __restrict__ raw,
                                                const int* __restrict__ perm, int Vsites, int V,
                                                size_t comp_stride, size_t par_stride, int Vh) {
  const int ds = blockIdx.x * LB + threadIdx.x;
  if (ds >= Vsites) return;
  const int i = perm[ds];
  const int e = i / V, lane = i - e * V;
  const int par = ds / Vh, idx = ds - par * Vh;
  const double* p = raw + (size_t)e * NC * 2 * V + lane;
#pragma unroll
  for (int k = 0; k < NC; k++) {
    double2 v;
    v.x = p[(size_t)(2 * k) * V];
    v.y = p[(size_t)(2 * k + 1) * V];
    dst[par * par_stride + k * comp_stride + idx] = v;
  }
}
template <int NC>
__global__ void __launch_bounds__(LB) k_scatter(double* __restrict__ raw, const double2* __restrict__ src,
                                                 const int* __restrict__ perm, int Vsites, int V,
                                                 size_t comp_stride, size_t par_stride, int Vh) {
  const int ds = blockIdx.x * LB + threadIdx.x;
  if (ds >= Vsites) return;
  const int i = perm[ds];
  const int e = i / V, lane = i - e * V;
  const int par = ds / Vh, idx = ds - par * Vh;
  double* p = raw + (size_t)e * NC * 2 * V + lane;
#pragma unroll
  for (int k = 0; k < NC; k++) {
    const double2 v = src[par * par_stride + k * comp_stride + idx];
    p[(size_t)(2 * k) * V] = v.x;
    p[(size_t)(2 * k + 1) * V] = v.y;
  }
}

int upload_vector(Field& f, const double* host) {
  Context& c = ctx();
  QB_CHECK(ensure_perm());
  const size_t bytes = sizeof(double) * 6 * (size_t)c.g.V;
  QB_CHECK(ensure_stage(bytes));
  QB_CUDA(cudaMemcpyAsync(c.d_stage, host, bytes, cudaMemcpyHostToDevice, c.stream));
  const int grid = (c.g.V + LB - 1) / LB;
  k_gather<3><<<grid, LB, 0, c.stream>>>(f.d, c.d_stage, c.d_perm, c.g.V, c.permV, (size_t)c.g.Vh,
                                         (size_t)3 * c.g.Vh, c.g.Vh);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

int download_vector(const Field& f, double* host) {
  Context& c = ctx();
  QB_CHECK(ensure_perm());
  const size_t bytes = sizeof(double) * 6 * (size_t)c.g.V;
  QB_CHECK(ensure_stage(bytes));
  const int grid = (c.g.V + LB - 1) / LB;
  k_scatter<3><<<grid, LB, 0, c.stream>>>(c.d_stage, f.d, c.d_perm, c.g.V, c.permV, (size_t)c.g.Vh,
                                          (size_t)3 * c.g.Vh, c.g.Vh);
  count_launch();
  QB_CUDA(cudaGetLastError());
  QB_CUDA(cudaMemcpyAsync(host, c.d_stage, bytes, cudaMemcpyDeviceToHost, c.stream));
  QB_CUDA(cudaStreamSynchronize(c.stream));
  return check_device_error();
}

int upload_links(Gauge& g, const double* const* links) {
  Context& c = ctx();
  QB_CHECK(ensure_perm());
  const size_t bytes = sizeof(double) * 18 * (size_t)c.g.V;
  QB_CHECK(ensure_stage(bytes));
  const int grid = (c.g.V + LB - 1) / LB;
  for (int k = 0; k < g.nlinks; k++) {
    if (!links[k]) return fail("gauge_load: links[%d] is NULL", k);
    QB_CUDA(cudaMemcpyAsync(c.d_stage, links[k], bytes, cudaMemcpyHostToDevice, c.stream));
    double2* dst = g.d + (size_t)k * 2 * 9 * c.g.Vh;
    k_gather<9><<<grid, LB, 0, c.stream>>>(dst, c.d_stage, c.d_perm, c.g.V, c.permV, (size_t)c.g.Vh,
                                           (size_t)9 * c.g.Vh, c.g.Vh);
    count_launch();
    QB_CUDA(cudaGetLastError());
  }
  QB_CUDA(cudaStreamSynchronize(c.stream));
  g.loaded = true;
  return compress_links(g);
}

// Device links back to native QEX host memory (the inverse of upload_links): what a caller needs
// after smearing on the device, e.g. to measure or to save the smeared links.
int download_links(const Gauge& g, double* const* links) {
  Context& c = ctx();
  QB_CHECK(ensure_perm());
  if (!g.loaded) return fail("gauge_download: gauge holds no links");
  const size_t bytes = sizeof(double) * 18 * (size_t)c.g.V;
  QB_CHECK(ensure_stage(bytes));
  const int grid = (c.g.V + LB - 1) / LB;
  for (int k = 0; k < g.nlinks; k++) {
    if (!links[k]) return fail("gauge_download: links[%d] is NULL", k);
    const double2* src = g.d + (size_t)k * 2 * 9 * c.g.Vh;
    k_scatter<9><<<grid, LB, 0, c.stream>>>(c.d_stage, src, c.d_perm, c.g.V, c.permV, (size_t)c.g.Vh,
                                            (size_t)9 * c.g.Vh, c.g.Vh);
    count_launch();
    QB_CUDA(cudaGetLastError());
    QB_CUDA(cudaMemcpyAsync(links[k], c.d_stage, bytes, cudaMemcpyDeviceToHost, c.stream));
    QB_CUDA(cudaStreamSynchronize(c.stream));
  }
  return 0;
}

// Staggered.rephase (stagD.nim:72-80): setBC (gaugeUtils.nim:124-131) then
// stagPhase with phases {8,9,11,0} (stagD.nim:509-522), on GLOBAL coordinates.
__global__ void __launch_bounds__(LB) k_rephase(double2* u, int Vh, int X0, int X1, int X2, int X3,
                                                 int o0, int o1, int o2, int o3, int T) {
  const int ds = blockIdx.x * LB + threadIdx.x;
  if (ds >= 2 * Vh) return;
  const int par = ds / Vh, idx = ds - par * Vh;
  const int X[4] = {X0, X1, X2, X3};
  const Site s = site_from_idx(idx, par, X);
  const int x0 = s.x[0] + o0, x1 = s.x[1] + o1, x3 = s.x[3] + o3;
  (void)o2;
  const int flip[4] = {x3 & 1, (x0 + x3) & 1, (x0 + x1 + x3) & 1, (x3 == T - 1) ? 1 : 0};
#pragma unroll
  for (int mu = 0; mu < 4; mu++) {
    if (flip[mu]) {
      double2* p = u + ((size_t)(mu * 2 + par) * 9) * Vh + idx;
#pragma unroll
      for (int e = 0; e < 9; e++) { double2 v = p[(size_t)e * Vh]; v.x = -v.x; v.y = -v.y; p[(size_t)e * Vh] = v; }
    }
  }
}

// ---- 12-real link storage -------------------------------------------------------------------
// A link class (one-hop / three-hop) whose matrices are s * W, W in SU(3), with one |s| for the
// whole field is also stored as rows 0 and 1 (96 B instead of 144 B per link); the sweeps rebuild
// row 2 = (sigma/|s|) conj(row0 x row1).  Naive links with staggered phases and antiperiodic BC
// are +-SU(3); Naik links are c * (product of three unitarised links).  Fat7-smeared links are not
// of this form and stay at 18 reals.  k_compress writes the rows, the sign bits and the evidence
// (max deviation of the rebuilt matrix, spread of |s|^2) the host decides on.
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// mode0 / mode1: storage tried for class 0 / 1: 12, 14 or 0 (class left alone).
// 14-real form: row2 = phi conj(row0 x row1) with phi = <c, row2> / <c, c> stored as the 7th element; the
// evidence is the deviation of the rebuilt third row alone (no common scale, no sign bit).
template <int NL>
__global__ void __launch_bounds__(LB) k_compress(const double2* __restrict__ u, double2* __restrict__ u12_0,
                                                  double2* __restrict__ u12_1, unsigned short* __restrict__ sgn,
                                                  unsigned long long* stats, int Vh, int mode0, int mode1) {
  const int ds = blockIdx.x * LB + threadIdx.x;
  const bool live = ds < 2 * Vh;
  const int par = live ? ds / Vh : 0, idx = live ? ds - par * Vh : 0;
  double dev[2] = {0.0, 0.0}, smin[2] = {1e300, 1e300}, smax[2] = {0.0, 0.0};
  unsigned int bits = 0;
  if (live) {
    bits = compress_site<NL>(u, u12_0, u12_1, Vh, par, idx, mode0, mode1, dev, smin, smax);
    if (sgn) sgn[ds] = (unsigned short)bits;
  }
#pragma unroll
  for (int j = 0; j < NL; j++) {
    const double a = warp_max(dev[j]), b = warp_min(smin[j]), cmax = warp_max(smax[j]);
    if ((threadIdx.x & 31) == 0) {   // non-negative doubles order like their bit patterns
      atomicMax(stats + j, (unsigned long long)__double_as_longlong(a));
      atomicMin(stats + 2 + j, (unsigned long long)__double_as_longlong(b));
      atomicMax(stats + 4 + j, (unsigned long long)__double_as_longlong(cmax));
    }
  }
}

// high byte of sgn: the sign bits of the links the BACKWARD hops read, U_{mu,j}(s - h_j mu)
template <int NL>
__global__ void __launch_bounds__(LB) k_sign_bwd(unsigned short* sgn, int Vh, int X0, int X1, int X2, int X3) {
  const int ds = blockIdx.x * LB + threadIdx.x;
  if (ds >= 2 * Vh) return;
  const int par = ds / Vh, idx = ds - par * Vh;
  const int X[4] = {X0, X1, X2, X3};
  unsigned int bits = sgn[ds] & 0xffu;
  bits = sign_bwd_site<NL>(sgn, Vh, par, idx, X, bits);
  sgn[ds] = (unsigned short)bits;
}

void free_compressed(Gauge& g) {
  for (int j = 0; j < 2; j++) {
    if (g.d12[j]) cudaFree(g.d12[j]);
    g.d12[j] = nullptr; g.recon[j] = 18; g.inv[j] = 1.0;
  }
  if (g.sgn) cudaFree(g.sgn);
  g.sgn = nullptr;
}

// Tolerances of the 12-real form: rebuilt matrix within 2e-14 |s| of the stored one, |s|^2 equal
// over the field to 1e-13 relative (the Dslash parity bar is 1e-13).  14-real form: the deviation bar alone.
constexpr double RECON_DEV_TOL = 2e-14;
constexpr double RECON_SCALE_TOL = 1e-13;

int compress_links(Gauge& g) {
  Context& c = ctx();
  QB_CUDA(cudaStreamSynchronize(c.stream));
  free_compressed(g);
  if ((!g.want12 && !g.want14) || !g.loaded) return 0;
  const int nl = g.nlinks / 4;
  int mode[2] = {0, 0};
  bool any12 = false, anym = false;
  for (int j = 0; j < nl; j++) {
    // each rank decides for its own block of the lattice: the halo layout does not depend on the storage form
    const bool w14 = ((g.want14 >> j) & 1) != 0;
    mode[j] = w14 ? 14 : (((g.want12 >> j) & 1) ? 12 : 0);
    if (!mode[j]) continue;
    anym = true; any12 = any12 || mode[j] == 12;
    const size_t n = (size_t)4 * 2 * (mode[j] == 14 ? 7 : 6) * c.g.Vh;
    QB_CUDA(cudaMalloc(&g.d12[j], sizeof(double2) * n));
  }
  if (!anym) return 0;
  if (any12) QB_CUDA(cudaMalloc(&g.sgn, sizeof(unsigned short) * 2 * (size_t)c.g.Vh));
  unsigned long long* stats = reinterpret_cast<unsigned long long*>(c.d_scal + 56);
  unsigned long long init[6] = {0, 0, 0x7ff0000000000000ull, 0x7ff0000000000000ull, 0, 0};
  QB_CUDA(cudaMemcpyAsync(stats, init, sizeof init, cudaMemcpyHostToDevice, c.stream));
  const int grid = (2 * c.g.Vh + LB - 1) / LB;
  if (nl == 2) k_compress<2><<<grid, LB, 0, c.stream>>>(g.d, g.d12[0], g.d12[1], g.sgn, stats, c.g.Vh, mode[0], mode[1]);
  else k_compress<1><<<grid, LB, 0, c.stream>>>(g.d, g.d12[0], nullptr, g.sgn, stats, c.g.Vh, mode[0], 0);
  count_launch();
  if (any12) {
    if (nl == 2) k_sign_bwd<2><<<grid, LB, 0, c.stream>>>(g.sgn, c.g.Vh, c.g.X[0], c.g.X[1], c.g.X[2], c.g.X[3]);
    else k_sign_bwd<1><<<grid, LB, 0, c.stream>>>(g.sgn, c.g.Vh, c.g.X[0], c.g.X[1], c.g.X[2], c.g.X[3]);
    count_launch();
  }
  QB_CUDA(cudaGetLastError());
  double hs[6];
  QB_CHECK(fetch_scalars(c.d_scal + 56, hs, 6));
  bool keep_sgn = false;
  for (int j = 0; j < nl; j++) {
    if (!mode[j]) continue;
    const double dev = hs[j], smin = hs[2 + j], smax = hs[4 + j];
    g.dev[j] = dev;
    bool ok = dev <= RECON_DEV_TOL && smin > 0.0;
    if (mode[j] == 12) ok = ok && (smax - smin) <= RECON_SCALE_TOL * smax;
    if (ok) {
      g.recon[j] = mode[j];
      if (mode[j] == 12) { g.inv[j] = 1.0 / sqrt(0.5 * (smin + smax)); keep_sgn = true; }
    } else {
      cudaFree(g.d12[j]); g.d12[j] = nullptr; g.recon[j] = 18;
    }
  }
  if (!keep_sgn && g.sgn) { cudaFree(g.sgn); g.sgn = nullptr; }
  return 0;
}

int rephase(Gauge& g) {
  Context& c = ctx();
  if (g.nlinks != 4) return fail("gauge_rephase: only one-hop link sets (nlinks=4) carry phases here");
  if (!g.loaded) return fail("gauge_rephase: gauge not loaded");
  const Geom& G = c.g;
  const int grid = (G.V + LB - 1) / LB;
  k_rephase<<<grid, LB, 0, c.stream>>>(g.d, G.Vh, G.X[0], G.X[1], G.X[2], G.X[3], G.rc[0] * G.X[0],
                                       G.rc[1] * G.X[1], G.rc[2] * G.X[2], G.rc[3] * G.X[3], G.lat[3]);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return compress_links(g);   // the signs changed: refresh the 12-real copy if one is wanted
}

}  // namespace qb
