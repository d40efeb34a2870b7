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
  return 0;
}

}  // namespace qb
