// Per-site random number field on the device: the input generator of the reference's benchmarks and tests
// (SURVEY.md 8d), so that lattices of any size get QEX's own synthetic fields without a host pass.
//
//   newRNGField(RngMilc6, lo, seed)   src/rng/distributionUtils.nim:306-331   one generator per site, seeded with
//                                                                             the GLOBAL lexicographic site index
//   RngMilc6 seedX / nextI / uniform  src/rng/milcrng.nim:92-154
//   gaussian (non-FUELCompat)         src/rng/milcrng.nim:158-188
//   v.gaussian r, g.random r          src/rng/distributionUtils.nim:65-97, src/gauge/gaugeUtils.nim:1348-1355,1424-1429
//   p.randomTAH r                     src/gauge/gaugeUtils.nim:1353-1383
//   projectSU                         src/maths/matrixFunctions.nim:359-370 (through project_u of linkops.cuh)
//
// The integer streams are bit-identical with QEX's.  A deviate is sqrt(-2 ln(u1 + 1e-307)) cos(2 pi u2) taken in
// double and rounded to float32 (the one reading of milcrng.nim:183-188 that reproduces the reference's golden
// plaquettes, tests/reprod/trandgauge.nim:17); the device's log/cos differ from glibc's in the last bit of the DOUBLE now and
// then, which survives the float32 rounding about once in 1e8 draws (tests/test_gpu_rng.py counts them).
#include "stencil.cuh"

namespace qb {

constexpr int RNG_BLOCK = 128;
constexpr unsigned int RNG_MASK = 0x00FFFFFFu;

struct Rng9 { unsigned int r0, r1, r2, r3, r4, r5, r6, ic, mult; };

__device__ __forceinline__ unsigned int rng_next(Rng9& p) {   // nextI, milcrng.nim:121-134
  const unsigned int t = (((p.r5 >> 7) | (p.r6 << 17)) ^ ((p.r4 >> 1) | (p.r5 << 23))) & RNG_MASK;
  p.r6 = p.r5; p.r5 = p.r4; p.r4 = p.r3; p.r3 = p.r2; p.r2 = p.r1; p.r1 = p.r0; p.r0 = t;
  const unsigned int s = p.ic * p.mult + 12345u;
  p.ic = s;
  return t ^ ((s >> 8) & RNG_MASK);
}
__device__ __forceinline__ float rng_uniform(Rng9& p) {        // uniform, milcrng.nim:150-154
  return (1.0f / 16777216.0f) * (float)rng_next(p);
}
__device__ __forceinline__ double rng_gaussian(Rng9& p) {      // gaussian, milcrng.nim:158-188
  const double v = (double)rng_uniform(p);
  const float u = rng_uniform(p);
  const double r = sqrt(-2.0 * log(v + 9.999999999999999e-308));
  const double ph = (double)u * 2.0 * 3.1415926535897932384626433;
  return (double)(float)(r * cos(ph));
}
__device__ __forceinline__ Rng9 rng_load(const unsigned int* st, size_t V, size_t i) {
  Rng9 p;
  p.r0 = st[i]; p.r1 = st[V + i]; p.r2 = st[2 * V + i]; p.r3 = st[3 * V + i]; p.r4 = st[4 * V + i];
  p.r5 = st[5 * V + i]; p.r6 = st[6 * V + i]; p.ic = st[7 * V + i]; p.mult = st[8 * V + i];
  return p;
}
__device__ __forceinline__ void rng_store(unsigned int* st, size_t V, size_t i, const Rng9& p) {
  st[i] = p.r0; st[V + i] = p.r1; st[2 * V + i] = p.r2; st[3 * V + i] = p.r3; st[4 * V + i] = p.r4;
  st[5 * V + i] = p.r5; st[6 * V + i] = p.r6; st[7 * V + i] = p.ic;
}

struct RngGeom { int X[4], off[4], lat[4], Vh; };

// global lexicographic index of device site (par, idx): x0 fastest (distributionUtils.nim:319-324)
__device__ __forceinline__ unsigned int global_lex(const RngGeom& G, int par, int idx) {
  const LSite s = lsite(idx, par, G.X);
  unsigned int li = (unsigned int)(s.x[3] + G.off[3]);
  li = li * G.lat[2] + (s.x[2] + G.off[2]);
  li = li * G.lat[1] + (s.x[1] + G.off[1]);
  li = li * G.lat[0] + (s.x[0] + G.off[0]);
  return li;
}

__global__ void __launch_bounds__(RNG_BLOCK) k_rng_seed(unsigned int* st, const RngGeom G, unsigned int seed0) {
  const int t = blockIdx.x * RNG_BLOCK + threadIdx.x;
  if (t >= 2 * G.Vh) return;
  const int par = t >= G.Vh, idx = t - par * G.Vh;
  const unsigned int index = global_lex(G, par, idx);
  const size_t V = 2 * (size_t)G.Vh;
  unsigned int seed = seed0;                                   // seedX, milcrng.nim:92-111
  const unsigned int a = 69607u + 8u * index;
#pragma unroll
  for (int k = 0; k < 7; k++) {
    seed = a * seed + 12345u;
    st[(size_t)k * V + t] = (seed >> 8) & RNG_MASK;
  }
  seed = a * seed + 12345u;
  st[7 * V + t] = seed;
  st[8 * V + t] = 100005u + 8u * index;
}

// v.gaussian r: six deviates per site, colour by colour, re then im
__global__ void __launch_bounds__(RNG_BLOCK) k_rng_vector(unsigned int* st, double2* v, int Vh) {
  const int t = blockIdx.x * RNG_BLOCK + threadIdx.x;
  if (t >= 2 * Vh) return;
  const int par = t >= Vh, idx = t - par * Vh;
  const size_t V = 2 * (size_t)Vh;
  Rng9 p = rng_load(st, V, t);
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const double re = rng_gaussian(p);
    const double im = rng_gaussian(p);
    v[((size_t)par * 3 + c) * Vh + idx] = make_double2(re, im);
  }
  rng_store(st, V, t, p);
}

// MODE 0: g[mu].gaussian r then projectSU (randomSU); MODE 1: randTah3 (8 deviates per site)
template <int MODE>
__global__ void __launch_bounds__(RNG_BLOCK) k_rng_matrix(unsigned int* st, double2* f, int Vh) {
  const int t = blockIdx.x * RNG_BLOCK + threadIdx.x;
  if (t >= 2 * Vh) return;
  const int par = t >= Vh, idx = t - par * Vh;
  const size_t V = 2 * (size_t)Vh;
  Rng9 p = rng_load(st, V, t);
  M3 m;
  if (MODE == 0) {
#pragma unroll
    for (int e = 0; e < 9; e++) {
      const double re = rng_gaussian(p);
      const double im = rng_gaussian(p);
      m.e[e] = make_double2(re, im);
    }
    m = project_u(m);
    // projectSU: multiply by det^(-1/3) (matrixFunctions.nim:359-370)
    const double2 a = c_mul(m.e[0], m.e[4]), b = c_mul(m.e[1], m.e[3]);
    const double2 c = c_mul(m.e[2], m.e[3]), d = c_mul(m.e[0], m.e[5]);
    const double2 g = c_mul(m.e[1], m.e[5]), h = c_mul(m.e[2], m.e[4]);
    const double2 t0 = c_mul(make_double2(a.x - b.x, a.y - b.y), m.e[8]);
    const double2 t1 = c_mul(make_double2(c.x - d.x, c.y - d.y), m.e[7]);
    const double2 t2 = c_mul(make_double2(g.x - h.x, g.y - h.y), m.e[6]);
    const double2 det = make_double2(t0.x + t1.x + t2.x, t0.y + t1.y + t2.y);
    const double ph = (1.0 / -3.0) * atan2(det.y, det.x);
    const double2 w = make_double2(cos(ph), sin(ph));
#pragma unroll
    for (int e = 0; e < 9; e++) m.e[e] = c_mul(w, m.e[e]);
  } else {
    const double s2 = 0.70710678118654752440, s3 = 0.57735026918962576450;
    const double r3 = s2 * rng_gaussian(p);
    const double r8 = s2 * s3 * rng_gaussian(p);
    const double r01 = s2 * rng_gaussian(p), r02 = s2 * rng_gaussian(p), r12 = s2 * rng_gaussian(p);
    const double i01 = s2 * rng_gaussian(p), i02 = s2 * rng_gaussian(p), i12 = s2 * rng_gaussian(p);
    m.e[0] = make_double2(0.0, r8 + r3); m.e[4] = make_double2(0.0, r8 - r3); m.e[8] = make_double2(0.0, -2 * r8);
    m.e[1] = make_double2(r01, i01); m.e[3] = make_double2(-r01, i01);
    m.e[2] = make_double2(r02, i02); m.e[6] = make_double2(-r02, i02);
    m.e[5] = make_double2(r12, i12); m.e[7] = make_double2(-r12, i12);
  }
  double2* o = f + (size_t)par * 9 * Vh + idx;
#pragma unroll
  for (int e = 0; e < 9; e++) o[(size_t)e * Vh] = m.e[e];
  rng_store(st, V, t, p);
}

__global__ void __launch_bounds__(RNG_BLOCK) k_unit_links(double2* f, int Vh, int nfields) {
  const size_t t = (size_t)blockIdx.x * RNG_BLOCK + threadIdx.x;
  const size_t n = (size_t)nfields * 2 * 9 * Vh;
  if (t >= n) return;
  const int e = (int)((t / Vh) % 9);
  f[t] = make_double2((e == 0 || e == 4 || e == 8) ? 1.0 : 0.0, 0.0);
}

int rng_create(unsigned int seed, RngField* r) {
  Context& c = ctx();
  const Geom& G = c.g;
  QB_CUDA(cudaMalloc(&r->st, sizeof(unsigned int) * 9 * (size_t)G.V));
  RngGeom g;
  for (int d = 0; d < 4; d++) { g.X[d] = G.X[d]; g.off[d] = G.rc[d] * G.X[d]; g.lat[d] = G.lat[d]; }
  g.Vh = G.Vh;
  k_rng_seed<<<(G.V + RNG_BLOCK - 1) / RNG_BLOCK, RNG_BLOCK, 0, c.stream>>>(r->st, g, seed);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}
void rng_destroy(RngField* r) {
  if (r->st) cudaFree(r->st);
  r->st = nullptr;
}
int rng_gaussian_vector(RngField& r, Field& f) {
  Context& c = ctx();
  k_rng_vector<<<(c.g.V + RNG_BLOCK - 1) / RNG_BLOCK, RNG_BLOCK, 0, c.stream>>>(r.st, f.d, c.g.Vh);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}
// mode 0: g.random r (field by field, like `for mu: randomSU(g[mu], r)`); mode 1: randomTAH
int rng_matrix_fields(RngField& r, Gauge& g, int mode) {
  Context& c = ctx();
  const int grid = (c.g.V + RNG_BLOCK - 1) / RNG_BLOCK;
  for (int k = 0; k < g.nlinks; k++) {
    double2* f = g.d + (size_t)k * 18 * c.g.Vh;
    if (mode == 0) k_rng_matrix<0><<<grid, RNG_BLOCK, 0, c.stream>>>(r.st, f, c.g.Vh);
    else k_rng_matrix<1><<<grid, RNG_BLOCK, 0, c.stream>>>(r.st, f, c.g.Vh);
    count_launch();
  }
  QB_CUDA(cudaGetLastError());
  g.loaded = true;
  return compress_links(g);
}
// ---- reunit (src/examples/staghmc_sh.nim:247-257): checkSU, projectSU on every link, checkSU again --------------
// checkSU of one matrix = |x^+ x - 1|^2 + |det x - 1|^2 (maths/matrixFunctions.nim:392-396)
__device__ __forceinline__ double2 m3_det(const M3& m) {
  const double2 a = c_mul(m.e[0], m.e[4]), b = c_mul(m.e[1], m.e[3]);
  const double2 c = c_mul(m.e[2], m.e[3]), d = c_mul(m.e[0], m.e[5]);
  const double2 g = c_mul(m.e[1], m.e[5]), h = c_mul(m.e[2], m.e[4]);
  const double2 t0 = c_mul(make_double2(a.x - b.x, a.y - b.y), m.e[8]);
  const double2 t1 = c_mul(make_double2(c.x - d.x, c.y - d.y), m.e[7]);
  const double2 t2 = c_mul(make_double2(g.x - h.x, g.y - h.y), m.e[6]);
  return make_double2(t0.x + t1.x + t2.x, t0.y + t1.y + t2.y);
}
__device__ __forceinline__ double check_su(const M3& m) {
  double d = 0.0;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      double2 s = make_double2(i == j ? -1.0 : 0.0, 0.0);   // (x^+ x)_ij = sum_k conj(x_ki) x_kj
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const double2 a = m.e[k * 3 + i], b = m.e[k * 3 + j];
        s.x += a.x * b.x + a.y * b.y;
        s.y += a.x * b.y - a.y * b.x;
      }
      d += s.x * s.x + s.y * s.y;
    }
  const double2 det = m3_det(m);
  return d + (det.x - 1.0) * (det.x - 1.0) + det.y * det.y;
}
// acc[0] sum before, [1] sum after (atomic adds: a diagnostic, like the echo of the reference), [2],[3] maxima as bits
__global__ void __launch_bounds__(RNG_BLOCK) k_reunit(double2* f, int Vh, int nfields, double* acc) {
  const size_t t = (size_t)blockIdx.x * RNG_BLOCK + threadIdx.x;
  const size_t n = (size_t)nfields * 2 * Vh;
  double db = 0.0, da = 0.0;
  if (t < n) {
    const size_t fp = t / Vh;            // field * 2 + parity
    const int idx = (int)(t - fp * Vh);
    double2* o = f + fp * 9 * Vh + idx;
    M3 m;
#pragma unroll
    for (int e = 0; e < 9; e++) m.e[e] = o[(size_t)e * Vh];
    db = check_su(m);
    m = project_u(m);                    // projectSU = projectU, then det^(-1/3) (matrixFunctions.nim:359-370)
    const double2 det = m3_det(m);
    const double ph = (1.0 / -3.0) * atan2(det.y, det.x);
    const double2 w = make_double2(cos(ph), sin(ph));
#pragma unroll
    for (int e = 0; e < 9; e++) m.e[e] = c_mul(w, m.e[e]);
    da = check_su(m);
#pragma unroll
    for (int e = 0; e < 9; e++) o[(size_t)e * Vh] = m.e[e];
  }
  double sb = db, sa = da, mb = db, ma = da;
#pragma unroll
  for (int o2 = 16; o2 > 0; o2 >>= 1) {
    sb += __shfl_down_sync(0xffffffffu, sb, o2); sa += __shfl_down_sync(0xffffffffu, sa, o2);
    mb = fmax(mb, __shfl_down_sync(0xffffffffu, mb, o2)); ma = fmax(ma, __shfl_down_sync(0xffffffffu, ma, o2));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(acc + 0, sb); atomicAdd(acc + 1, sa);
    atomicMax(reinterpret_cast<unsigned long long*>(acc + 2), (unsigned long long)__double_as_longlong(mb));
    atomicMax(reinterpret_cast<unsigned long long*>(acc + 3), (unsigned long long)__double_as_longlong(ma));
  }
}
__global__ void k_xsum4(XRank* xr, double* acc) {   // sums [0],[1] and maxima [2],[3] over the ranks of the box
  const double a = xrank_sum(xr, acc[0]), b = xrank_sum(xr, acc[1]);
  const double c = xrank_sum<true>(xr, acc[2]), d = xrank_sum<true>(xr, acc[3]);
  if ((threadIdx.x & 31) == 0) { acc[0] = a; acc[1] = b; acc[2] = c; acc[3] = d; }
}
// out[0..3] = avg, max before; avg, max after -- normalised like checkSU (gauge/gaugeUtils.nim:1407-1422)
int reunit_links(Gauge& g, double out[4]) {
  Context& c = ctx();
  if (!g.loaded) return fail("gauge_reunit: gauge not loaded");
  double* acc = c.d_scal + 40;
  QB_CUDA(cudaMemsetAsync(acc, 0, sizeof(double) * 4, c.stream));
  const size_t n = (size_t)g.nlinks * 2 * c.g.Vh;
  k_reunit<<<(unsigned)((n + RNG_BLOCK - 1) / RNG_BLOCK), RNG_BLOCK, 0, c.stream>>>(g.d, c.g.Vh, g.nlinks, acc);
  count_launch();
  if (c.g.nranks > 1) {
    if (c.p2p) {
      k_xsum4<<<1, 32, 0, c.stream>>>(c.d_xr, acc);
      count_launch();
    } else if (c.nccl) {
      QB_NCCL(nccl().AllReduce(acc, acc, 2, ncclDouble, ncclSum, c.nccl, c.stream));
      QB_NCCL(nccl().AllReduce(acc + 2, acc + 2, 2, ncclDouble, ncclMax, c.nccl, c.stream));
    }
  }
  QB_CUDA(cudaGetLastError());
  double h[4];
  QB_CHECK(fetch_scalars(acc, h, 4));
  const double cn = 2.0 * (3 * 3 + 1);
  const double vol = (double)c.g.V * c.g.nranks;
  if (out) {
    out[0] = sqrt(h[0] / (cn * g.nlinks * vol)); out[1] = sqrt(h[2] / cn);
    out[2] = sqrt(h[1] / (cn * g.nlinks * vol)); out[3] = sqrt(h[3] / cn);
  }
  return compress_links(g);
}

int unit_links(Gauge& g) {
  Context& c = ctx();
  const size_t n = (size_t)g.nlinks * 18 * c.g.Vh;
  k_unit_links<<<(unsigned)((n + RNG_BLOCK - 1) / RNG_BLOCK), RNG_BLOCK, 0, c.stream>>>(g.d, c.g.Vh, g.nlinks);
  count_launch();
  QB_CUDA(cudaGetLastError());
  g.loaded = true;
  return compress_links(g);
}

}  // namespace qb
