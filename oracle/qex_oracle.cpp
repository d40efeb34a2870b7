// qex_oracle.cpp -- CPU ORACLE (TEST INFRASTRUCTURE ONLY, NOT PRODUCT CODE)
//
// A plain C++17/OpenMP restatement of the arithmetic of QEX's staggered
// Dslash + CG hot path.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load this library, and
// only as the checker or the timed CPU baseline.  The product
// (qex_b200/csrc) never links, imports or calls anything in oracle/.
//
// Parity pins (numbers the reference itself holds):
//  (1) inputs -- RngMilc6 seeding + Box-Muller gaussian + projectSU + plaq against the known-answer test
//      tests/reprod/trandgauge.nim:14-17 (six plaquettes of g.random on 8^4, default seed 17^7);
//  (2) operator and solver -- the first "Begin H" line of the golden logs tests/extra/staghmc_sh/ref.{0,1,2}
//      (Sf per pseudofermion and T, the reference's own bar rel 2e-11, tests/extra/staghmc_sh/run:43), which is
//      computed before any MD step from randomTAH, setBC/stagPhase, D with negative masses, `solve` at
//      r2req = 1e-24 with its even-odd reconstruction and norm2: reproduced to a few ulp, iteration counts held
//      to the Solver[action] statistics of the same logs -- tests/test_oracle_golden.py
//      (test_trandgauge_golden_plaquettes, test_staghmc_sh_begin_H_golden, test_staghmc_sh_pin_discriminates).
// The reference publishes no golden Dslash vector on a random gauge; there the restatement is
// additionally held by analytic identities (free-field eigenvalues, anti-Hermiticity, D^+D block identity,
// solve(Dx)=x).  "End H" of the golden logs needs the MD integrator of an external package and is not reproduced.
// The reference itself (Nim + MPI + QMP) cannot be built in this image.
//
// Field memory layout used by the oracle == QEX's own layout for a single rank
// with V=1 (innerGeom = 1,1,1,1): site index i = lex(x)/2 (+ nSites/2 for odd
// sites), lex with x[0] fastest (layout/qlayout.nim:110-131); a colour vector
// is double[3][2] per site, a colour matrix double[3][3][2] row-major
// (maths/matrixConcept.nim:117-123, maths/complexType.nim:11-13).
//
// All file:line citations are relative to /root/reference.

#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <complex>
#include <omp.h>

typedef std::complex<double> cplx;

extern "C" {

// ---------------------------------------------------------------------------
// Layout index maps  (src/layout/qlayout.nim:10-185)
// ---------------------------------------------------------------------------
struct OrcLayout {
  int nd;
  int phys[4], rankg[4], inner[4], local[4], outer[4];
  int nSites, nEven, nOdd, nSitesOuter, nEvenOuter, nOddOuter, nSitesInner;
  int innerCb, innerCbDir;
};

// layoutSetupQ, qlayout.nim:10-64.  returns 0 on success.
int orc_layout_setup(OrcLayout* l, const int* phys, const int* rankg, const int* inner) {
  l->nd = 4;
  long pvol = 1, lvol = 1, ovol = 1;
  int icb = 0, icbd = -1;
  for (int i = 0; i < 4; i++) {
    l->phys[i] = phys[i]; l->rankg[i] = rankg[i]; l->inner[i] = inner[i];
    l->local[i] = phys[i] / rankg[i];
    l->outer[i] = l->local[i] / inner[i];
    pvol *= phys[i]; lvol *= l->local[i]; ovol *= l->outer[i];
    if (inner[i] > 1 && (l->outer[i] & 1) == 1) icb++;
    if (inner[i] == 1 && (l->outer[i] & 1) == 0) icbd = i;
  }
  if (icb == 0) icbd = 0;
  else {
    if (icbd < 0) return -1;
    icb = l->outer[icbd] / 2;
    if ((icb & 1) == 0) return -2;
  }
  l->nSites = (int)lvol;
  l->nOdd = (int)(lvol / 2);
  l->nEven = (int)(lvol - l->nOdd);
  l->nSitesOuter = (int)ovol;
  l->nOddOuter = (int)(ovol / 2);
  l->nEvenOuter = (int)(ovol - l->nOddOuter);
  l->nSitesInner = l->nSites / l->nSitesOuter;
  l->innerCb = icb;
  l->innerCbDir = icbd;
  return 0;
}

// layoutIndexQ, qlayout.nim:110-131.  coords are GLOBAL lattice coordinates.
void orc_layout_index(const OrcLayout* l, const int* coords_in, int* rank, int* index) {
  int c[4];
  for (int i = 0; i < 4; i++) c[i] = coords_in[i];
  int ri = 0;   // lexr_i: x[0] slowest (qlayout.nim:94-104)
  for (int i = 0; i < 4; i++) ri = ri * l->rankg[i] + ((c[i] / l->local[i]) % l->rankg[i]);
  int ii = 0;   // lex_i: x[0] fastest (qlayout.nim:80-91)
  for (int i = 3; i >= 0; i--) ii = ii * l->inner[i] + ((c[i] / l->outer[i]) % l->inner[i]);
  int ib = 0;
  for (int i = 0; i < 4; i++) {
    int xi = c[i] / l->outer[i];
    int xli = xi % l->inner[i];
    ib += xli * l->outer[i];
  }
  ib &= 1;
  c[l->innerCbDir] += l->innerCb * ib;
  int oi = 0;
  for (int i = 3; i >= 0; i--) oi = oi * l->outer[i] + (c[i] % l->outer[i]);
  c[l->innerCbDir] -= l->innerCb * ib;
  int p = 0;
  for (int i = 0; i < 4; i++) p += c[i];
  int oi2 = oi / 2;
  if (p & 1) oi2 = (oi + l->nSitesOuter) / 2;
  *rank = ri;
  *index = oi2 * l->nSitesInner + ii;
}

// layoutCoordQ, qlayout.nim:133-185 (inverse map), restated as a search-free
// closed form and checked against orc_layout_index by the tests.
void orc_layout_coord(const OrcLayout* l, int rank, int index, int* coords) {
  int cr[4];
  { int r = rank; for (int i = 3; i >= 0; i--) { cr[i] = r % l->rankg[i]; r /= l->rankg[i]; } }
  int p = 0;
  int ll = index % l->nSitesInner;
  int ib = 0;
  for (int i = 0; i < 4; i++) {
    int w = l->inner[i], wl = l->outer[i];
    int k = ll % w;
    int c = l->local[i] * cr[i] + k * wl;
    cr[i] = c; p += c; ll /= w; ib += k * wl;
  }
  ib &= 1;
  int ii = index / l->nSitesInner;
  if (ii >= l->nEvenOuter) { ii -= l->nEvenOuter; p++; }
  ii *= 2;
  for (int i = 0; i < 4; i++) {
    int wl = l->outer[i];
    int k = ii % wl;
    if (i == l->innerCbDir) k = (k + l->innerCb * ib) % wl;
    coords[i] = k; p += k; ii /= wl;
  }
  if (p & 1) {
    for (int i = 0; i < 4; i++) {
      int wl = l->outer[i];
      if (i == l->innerCbDir) coords[i] = (coords[i] + l->innerCb * ib) % wl;
      coords[i]++;
      if (coords[i] >= wl) {
        coords[i] = 0;
        if (i == l->innerCbDir) coords[i] = (coords[i] + l->innerCb * ib) % wl;
      } else {
        if (i == l->innerCbDir) coords[i] = (coords[i] + l->innerCb * ib) % wl;
        break;
      }
    }
  }
  for (int i = 0; i < 4; i++) coords[i] += cr[i];
}

// For every site of the global lattice, in the oracle's own (V=1, single rank)
// order: owning rank and QEX single-site index i = outer*V + lane in layout `l`.
void orc_layout_table(const OrcLayout* l, int* rank_out, int* index_out) {
  const int* L = l->phys;
  int vol = L[0] * L[1] * L[2] * L[3];
  int x[4];
  for (x[3] = 0; x[3] < L[3]; x[3]++)
  for (x[2] = 0; x[2] < L[2]; x[2]++)
  for (x[1] = 0; x[1] < L[1]; x[1]++)
  for (x[0] = 0; x[0] < L[0]; x[0]++) {
    int lex = ((x[3] * L[2] + x[2]) * L[1] + x[1]) * L[0] + x[0];
    int p = (x[0] + x[1] + x[2] + x[3]) & 1;
    int i = p ? (lex + vol) / 2 : lex / 2;
    orc_layout_index(l, x, &rank_out[i], &index_out[i]);
  }
}

// ---------------------------------------------------------------------------
// The oracle's working lattice: single rank, V=1.
// ---------------------------------------------------------------------------
struct Lat {
  int L[4];
  int vol, half;
  std::vector<int> fwd[4][2];  // fwd[mu][h==1?0:1][i] = index of x + h*mu
  std::vector<int> bwd[4][2];
  std::vector<short> x[4];     // coords of site i
};

static inline int lat_index(const int* L, const int* x) {
  int lex = ((x[3] * L[2] + x[2]) * L[1] + x[1]) * L[0] + x[0];
  int p = (x[0] + x[1] + x[2] + x[3]) & 1;
  int vol = L[0] * L[1] * L[2] * L[3];
  return p ? (lex + vol) / 2 : lex / 2;
}

void* orc_lat_new(const int* L) {
  Lat* l = new Lat;
  for (int i = 0; i < 4; i++) l->L[i] = L[i];
  l->vol = L[0] * L[1] * L[2] * L[3];
  l->half = l->vol / 2;
  for (int mu = 0; mu < 4; mu++) {
    l->x[mu].resize(l->vol);
    for (int h = 0; h < 2; h++) { l->fwd[mu][h].resize(l->vol); l->bwd[mu][h].resize(l->vol); }
  }
  int x[4];
  for (x[3] = 0; x[3] < L[3]; x[3]++)
  for (x[2] = 0; x[2] < L[2]; x[2]++)
  for (x[1] = 0; x[1] < L[1]; x[1]++)
  for (x[0] = 0; x[0] < L[0]; x[0]++) {
    int i = lat_index(L, x);
    for (int mu = 0; mu < 4; mu++) l->x[mu][i] = (short)x[mu];
    for (int mu = 0; mu < 4; mu++) for (int h = 0; h < 2; h++) {
      int hop = h ? 3 : 1;
      int y[4] = {x[0], x[1], x[2], x[3]};
      y[mu] = (x[mu] + hop) % L[mu];
      l->fwd[mu][h][i] = lat_index(L, y);
      y[mu] = ((x[mu] - hop) % L[mu] + L[mu]) % L[mu];
      l->bwd[mu][h][i] = lat_index(L, y);
    }
  }
  return l;
}
void orc_lat_free(void* p) { delete (Lat*)p; }
int orc_lat_vol(void* p) { return ((Lat*)p)->vol; }
void orc_lat_coords(void* p, short* out) {  // out[4][vol]
  Lat* l = (Lat*)p;
  for (int mu = 0; mu < 4; mu++) memcpy(out + (size_t)mu * l->vol, l->x[mu].data(), sizeof(short) * l->vol);
}

// ---------------------------------------------------------------------------
// RngMilc6  (src/rng/milcrng.nim:12-188, non-FUELCompat branch)
// ---------------------------------------------------------------------------
struct Rng { uint32_t r0, r1, r2, r3, r4, r5, r6, ic, mult; };

static const uint32_t INDX1 = 69607u, INDX2 = 8u, ADDEND = 12345u, MASK = 0x00FFFFFFu;

// seedX, milcrng.nim:92-111
static void rng_seed(Rng* p, uint32_t seed0, uint32_t index) {
  uint32_t seed = seed0;
  uint32_t* r[7] = {&p->r0, &p->r1, &p->r2, &p->r3, &p->r4, &p->r5, &p->r6};
  for (int k = 0; k < 7; k++) {
    seed = (INDX1 + INDX2 * index) * seed + ADDEND;
    *r[k] = (seed >> 8) & MASK;
  }
  seed = (INDX1 + INDX2 * index) * seed + ADDEND;
  p->ic = seed;
  p->mult = 100005u + 8u * index;
}
// nextI, milcrng.nim:121-134
static inline uint32_t rng_next(Rng* p) {
  uint32_t t = (((p->r5 >> 7) | (p->r6 << 17)) ^ ((p->r4 >> 1) | (p->r5 << 23))) & MASK;
  p->r6 = p->r5; p->r5 = p->r4; p->r4 = p->r3; p->r3 = p->r2; p->r2 = p->r1; p->r1 = p->r0; p->r0 = t;
  uint32_t s = p->ic * p->mult + ADDEND;
  p->ic = s;
  return t ^ ((s >> 8) & MASK);
}
// uniform, milcrng.nim:150-154: float32 result
static inline float rng_uniform(Rng* p) {
  const float SCALE = 1.0f / (float)0x01000000;
  return SCALE * (float)rng_next(p);
}
// gaussian, milcrng.nim:158-188 (non-FUELCompat): Box-Muller, one deviate per
// two uniforms.  The source mixes float32 uniforms with float64 maths
// (`var p = prn.uniform * 2.0 * PI`), and which products Nim takes in which
// precision is not obvious from the text.  It was settled EMPIRICALLY against the
// reference's known-answer plaquettes (tests/reprod/trandgauge.nim:14-17): of 18
// candidate readings exactly one reproduces all six to sum-diff^2 = 1.4e-31
// (bar: 1e-30): radius and phase in float64, the DEVIATE rounded to float32.
// gauss_variant 0 keeps the all-float64 reading (misses the golden at 1e-20) so
// the test can show the pin discriminates.
static int gauss_variant = 8;
void orc_set_gauss_variant(int v) { gauss_variant = v; }
int orc_get_gauss_variant() { return gauss_variant; }
static inline double rng_gaussian(Rng* p) {
  const double TINY = 9.999999999999999e-308;
  const double PI = 3.1415926535897932384626433;
  double v = (double)rng_uniform(p);
  float u = rng_uniform(p);
  double r = sqrt(-2.0 * log(v + TINY));
  double ph = (double)u * 2.0 * PI;
  double g = r * cos(ph);
  return gauss_variant == 0 ? g : (double)(float)g;
}

// newRNGField, distributionUtils.nim:306-331: one generator per site, seeded
// with the lexicographic index (x[0] fastest, x[nd-1] slowest) of its coords.
void* orc_rng_new(void* lat, uint64_t seed) {
  Lat* l = (Lat*)lat;
  Rng* r = (Rng*)malloc(sizeof(Rng) * l->vol);
  for (int i = 0; i < l->vol; i++) {
    long li = l->x[3][i];
    for (int d = 2; d >= 0; d--) li = li * l->L[d] + l->x[d][i];
    rng_seed(&r[i], (uint32_t)seed, (uint32_t)li);
  }
  return r;
}
void orc_rng_free(void* r) { free(r); }
uint32_t orc_rng_next_site(void* r, int site) { return rng_next(&((Rng*)r)[site]); }

// gaussian(Field, RNGField): distributionUtils.nim:13-24,65-97 -- every real
// component of the site object in row-major order, re then im.
void orc_gaussian(void* lat, void* rng, double* f, int nreal) {
  Lat* l = (Lat*)lat; Rng* r = (Rng*)rng;
  for (int i = 0; i < l->vol; i++)
    for (int k = 0; k < nreal; k++) f[(size_t)i * nreal + k] = rng_gaussian(&r[i]);
}

// randomTAH(Field, RNGField) for nc = 3: randTah3, gauge/gaugeUtils.nim:1353-1383 -- eight deviates per
// site in the order r3, r8, r01, r02, r12, i01, i02, i12 (each the float32 result of gaussian widened to
// float64, times sqrt(1/2)); the matrix is traceless anti-Hermitian with an imaginary diagonal.
void orc_random_tah(void* lat, void* rng, double* f) {
  Lat* l = (Lat*)lat; Rng* r = (Rng*)rng;
  const double s2 = 0.70710678118654752440, s3 = 0.57735026918962576450;
  for (int i = 0; i < l->vol; i++) {
    double* m = f + (size_t)i * 18;
    Rng* s = &r[i];
    double r3 = s2 * rng_gaussian(s);
    double r8 = s2 * s3 * rng_gaussian(s);
    double r01 = s2 * rng_gaussian(s), r02 = s2 * rng_gaussian(s), r12 = s2 * rng_gaussian(s);
    double i01 = s2 * rng_gaussian(s), i02 = s2 * rng_gaussian(s), i12 = s2 * rng_gaussian(s);
#define TAH_SET(a, b, re, im) do { m[((a) * 3 + (b)) * 2] = (re); m[((a) * 3 + (b)) * 2 + 1] = (im); } while (0)
    TAH_SET(0, 0, 0.0, r8 + r3); TAH_SET(1, 1, 0.0, r8 - r3); TAH_SET(2, 2, 0.0, -2 * r8);
    TAH_SET(0, 1, r01, i01); TAH_SET(1, 0, -r01, i01);
    TAH_SET(0, 2, r02, i02); TAH_SET(2, 0, -r02, i02);
    TAH_SET(1, 2, r12, i12); TAH_SET(2, 1, -r12, i12);
#undef TAH_SET
  }
}

// ---------------------------------------------------------------------------
// 3x3 complex algebra
// ---------------------------------------------------------------------------
struct M3 { cplx e[3][3]; };
struct V3 { cplx e[3]; };

static inline M3 m3_load(const double* p) {
  M3 m; for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) m.e[a][b] = cplx(p[(a * 3 + b) * 2], p[(a * 3 + b) * 2 + 1]); return m;
}
static inline void m3_store(double* p, const M3& m) {
  for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) { p[(a * 3 + b) * 2] = m.e[a][b].real(); p[(a * 3 + b) * 2 + 1] = m.e[a][b].imag(); }
}
static inline cplx cmul(cplx a, cplx b) {  // plain complex product (no inf/nan fix-ups)
  return cplx(a.real() * b.real() - a.imag() * b.imag(), a.real() * b.imag() + a.imag() * b.real());
}
static inline M3 m3_mul(const M3& a, const M3& b) {
  M3 r; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { cplx s = 0; for (int k = 0; k < 3; k++) s += cmul(a.e[i][k], b.e[k][j]); r.e[i][j] = s; } return r;
}
static inline M3 m3_adj(const M3& a) {
  M3 r; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.e[i][j] = std::conj(a.e[j][i]); return r;
}
// determinant, matrixFunctions.nim:66-77
static inline cplx m3_det(const M3& x) {
  return cmul(cmul(x.e[0][0], x.e[1][1]) - cmul(x.e[0][1], x.e[1][0]), x.e[2][2]) +
         cmul(cmul(x.e[0][2], x.e[1][0]) - cmul(x.e[0][0], x.e[1][2]), x.e[2][1]) +
         cmul(cmul(x.e[0][1], x.e[1][2]) - cmul(x.e[0][2], x.e[1][1]), x.e[2][0]);
}
// eigs3, matrixFunctions.nim:79-111
static inline void eigs3(double& e0, double& e1, double& e2, double tr, double p2, double det) {
  double tr3 = (1.0 / 3.0) * tr;
  double p23 = (1.0 / 3.0) * p2;
  double tr32 = tr3 * tr3;
  double q = fabs(0.5 * (p23 - tr32));
  double r = 0.25 * tr3 * (5 * tr32 - p2) - 0.5 * det;
  double sq = sqrt(q);
  double sq3 = q * sq;
  double isq3 = 1.0 / sq3;
  double isq3c = fmin(3e38, fmax(-3e38, isq3));
  double rsq3c = r * isq3c;
  double rsq3 = fmin(1.0, fmax(-1.0, rsq3c));
  double t = (1.0 / 3.0) * acos(rsq3);
  double st = sin(t), ct = cos(t);
  double sqc = sq * ct;
  double sqs = 1.73205080756887729352 * sq * st;
  double ll = tr3 + sqc;
  e0 = tr3 - 2 * sqc; e1 = ll + sqs; e2 = ll - sqs;
}
// rsqrtPHM3f + rsqrtPHM3, matrixFunctions.nim:126-181
static inline M3 rsqrtPH3(const M3& x) {
  double tr = (x.e[0][0] + x.e[1][1] + x.e[2][2]).real();
  M3 x2 = m3_mul(x, x);
  double p2 = (x2.e[0][0] + x2.e[1][1] + x2.e[2][2]).real();
  double det = m3_det(x).real();
  double l0, l1, l2;
  eigs3(l0, l1, l2, tr, p2, det);
  double sl0 = sqrt(fabs(l0)), sl1 = sqrt(fabs(l1)), sl2 = sqrt(fabs(l2));
  double u = sl0 + sl1 + sl2;
  double w = sl0 * sl1 * sl2;
  double d = w * (sl0 + sl1) * (sl0 + sl2) * (sl1 + sl2);
  double di = 1 / d;
  double c0 = (w * u * u + l0 * sl0 * (l1 + l2) + l1 * sl1 * (l0 + l2) + l2 * sl2 * (l0 + l1)) * di;
  double c1 = -(tr * u + w) * di;
  double c2 = u * di;
  M3 r;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    r.e[i][j] = c1 * x.e[i][j] + c2 * x2.e[i][j];
    if (i == j) r.e[i][j] += c0;
  }
  return r;
}
// projectU, matrixFunctions.nim:301-312: x (x^+ x + eps)^{-1/2}
static inline M3 projectU(const M3& x) {
  M3 t = m3_mul(m3_adj(x), x);
  for (int i = 0; i < 3; i++) t.e[i][i] += 1e-20;
  return m3_mul(x, rsqrtPH3(t));
}
// projectSU, matrixFunctions.nim:359-370
static inline M3 projectSU(const M3& x) {
  M3 m = projectU(x);
  cplx d = m3_det(m);
  double p = (1.0 / -3.0) * atan2(d.imag(), d.real());
  cplx ph(cos(p), sin(p));
  M3 r; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.e[i][j] = cmul(ph, m.e[i][j]);
  return r;
}
void orc_projectU(const double* in, double* out) { m3_store(out, projectU(m3_load(in))); }
void orc_projectSU(const double* in, double* out) { m3_store(out, projectSU(m3_load(in))); }

// random(g, r): gaugeUtils.nim:1424-1429 -> randomSU :1352-1354, mu by mu on the
// continuing per-site streams.  g = nmu contiguous fields [mu][vol][18].
void orc_random_gauge(void* lat, void* rng, double* g, int nmu) {
  Lat* l = (Lat*)lat;
  for (int mu = 0; mu < nmu; mu++) {
    double* gm = g + (size_t)mu * l->vol * 18;
    orc_gaussian(lat, rng, gm, 18);
#pragma omp parallel for
    for (int i = 0; i < l->vol; i++) m3_store(gm + (size_t)i * 18, projectSU(m3_load(gm + (size_t)i * 18)));
  }
}

// plaq, gaugeUtils.nim:213-284: pl[ip(mu,nu)] = sum_x Re tr[(U_mu(x)U_nu(x+mu))^+ (U_nu(x)U_mu(x+nu))] / (vol*6*3)
void orc_plaq(void* lat, const double* g, double* pl) {
  Lat* l = (Lat*)lat;
  size_t fs = (size_t)l->vol * 18;
  for (int mu = 1; mu < 4; mu++) for (int nu = 0; nu < mu; nu++) {
    double s = 0;
#pragma omp parallel for reduction(+ : s)
    for (int i = 0; i < l->vol; i++) {
      M3 unumu = m3_mul(m3_load(g + nu * fs + (size_t)i * 18), m3_load(g + mu * fs + (size_t)l->fwd[nu][0][i] * 18));
      M3 umunu = m3_mul(m3_load(g + mu * fs + (size_t)i * 18), m3_load(g + nu * fs + (size_t)l->fwd[mu][0][i] * 18));
      double d = 0;
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++)
        d += umunu.e[a][b].real() * unumu.e[a][b].real() + umunu.e[a][b].imag() * unumu.e[a][b].imag();
      s += d;
    }
    pl[(mu * (mu - 1)) / 2 + nu] = s / ((double)l->vol * 6.0 * 3.0);
  }
}

// setBC, gaugeUtils.nim:124-131: U_3 *= -1 on the last time slice
void orc_setbc(void* lat, double* g) {
  Lat* l = (Lat*)lat;
  double* gt = g + (size_t)3 * l->vol * 18;
  for (int i = 0; i < l->vol; i++)
    if (l->x[3][i] == l->L[3] - 1) for (int k = 0; k < 18; k++) gt[(size_t)i * 18 + k] *= -1;
}
// stagPhase, stagD.nim:509-522 with phases [8,9,11,0]
void orc_stagphase(void* lat, double* g) {
  Lat* l = (Lat*)lat;
  const int phases[4] = {8, 9, 11, 0};
  for (int mu = 0; mu < 4; mu++) {
    double* gm = g + (size_t)mu * l->vol * 18;
    for (int i = 0; i < l->vol; i++) {
      int s = 0;
      for (int k = 0; k < 4; k++) s += (phases[mu] >> k) & l->x[k][i];
      if (s & 1) for (int k = 0; k < 18; k++) gm[(size_t)i * 18 + k] *= -1;
    }
  }
}

// ---------------------------------------------------------------------------
// Staggered operator  (src/physics/stagD.nim)
// g = nd contiguous link fields [k][vol][18]; nd==4: hop 1 in dir k;
// nd==8: g[2mu] hop 1, g[2mu+1] hop 3 in dir mu  (stagD.nim:36-49,552-564).
// ---------------------------------------------------------------------------
static inline void mv_acc(double* acc, const double* u, const double* v, double sgn) {
  // acc += sgn * U v
  for (int a = 0; a < 3; a++) {
    double re = 0, im = 0;
    for (int b = 0; b < 3; b++) {
      double ur = u[(a * 3 + b) * 2], ui = u[(a * 3 + b) * 2 + 1];
      re += ur * v[2 * b] - ui * v[2 * b + 1];
      im += ur * v[2 * b + 1] + ui * v[2 * b];
    }
    acc[2 * a] += sgn * re; acc[2 * a + 1] += sgn * im;
  }
}
static inline void mv_adj_acc(double* acc, const double* u, const double* v, double sgn) {
  // acc += sgn * U^+ v
  for (int a = 0; a < 3; a++) {
    double re = 0, im = 0;
    for (int b = 0; b < 3; b++) {
      double ur = u[(b * 3 + a) * 2], ui = -u[(b * 3 + a) * 2 + 1];
      re += ur * v[2 * b] - ui * v[2 * b + 1];
      im += ur * v[2 * b + 1] + ui * v[2 * b];
    }
    acc[2 * a] += sgn * re; acc[2 * a + 1] += sgn * im;
  }
}

// stagD2 on one parity, stagD.nim:348-395:
//   r[s] = a*r[s] + b*x[s] + sum_k [ g_k(s) x(s+h_k) - g_k(s-h_k)^+ x(s-h_k) ]
// sgn=-1 gives the stagDM flavour (stagD.nim:278-313).  parity 0=even, 1=odd.
static void stag_sweep(Lat* l, int nd, const double* g, double* r, const double* x,
                       double a, const double* bx, double b, int parity, double sgn) {
  size_t fs = (size_t)l->vol * 18;
  int lo = parity ? l->half : 0, hi = parity ? l->vol : l->half;
#pragma omp parallel for schedule(static)
  for (int i = lo; i < hi; i++) {
    double acc[6] = {0, 0, 0, 0, 0, 0};
    for (int k = 0; k < nd; k++) {
      int mu = (nd == 8) ? k / 2 : k;
      int h = (nd == 8) ? (k & 1) : 0;
      int jf = l->fwd[mu][h][i], jb = l->bwd[mu][h][i];
      mv_acc(acc, g + k * fs + (size_t)i * 18, x + (size_t)jf * 6, 1.0);
      mv_adj_acc(acc, g + k * fs + (size_t)jb * 18, x + (size_t)jb * 6, -1.0);
    }
    for (int c = 0; c < 6; c++) {
      double ini = 0;
      if (a != 0.0) ini += a * r[(size_t)i * 6 + c];
      if (b != 0.0) ini += b * bx[(size_t)i * 6 + c];
      r[(size_t)i * 6 + c] = ini + sgn * acc[c];
    }
  }
}

void orc_stagD2(void* lat, int nd, const double* g, double* r, const double* x, double a, double b, int parity) {
  stag_sweep((Lat*)lat, nd, g, r, x, a, x, b, parity, 1.0);
}
// stagD, stagD.nim:405-409
void orc_stagD(void* lat, int nd, const double* g, double* r, const double* x, double m, double sc, double a, int parity) {
  Lat* l = (Lat*)lat;
  orc_stagD2(lat, nd, g, r, x, a / (0.5 * sc), m / (0.5 * sc), parity);
  int lo = parity ? l->half : 0, hi = parity ? l->vol : l->half;
#pragma omp parallel for
  for (int i = lo * 6; i < hi * 6; i++) r[i] = (0.5 * sc) * r[i];
}
// D / Ddag, stagD.nim:566-571
void orc_D(void* lat, int nd, const double* g, double* r, const double* x, double m) {
  orc_stagD(lat, nd, g, r, x, m, 1.0, 0.0, 0); orc_stagD(lat, nd, g, r, x, m, 1.0, 0.0, 1);
}
void orc_Ddag(void* lat, int nd, const double* g, double* r, const double* x, double m) {
  orc_stagD(lat, nd, g, r, x, m, -1.0, 0.0, 0); orc_stagD(lat, nd, g, r, x, m, -1.0, 0.0, 1);
}
// stagD2xx, stagD.nim:433-469: r[px] = 4 m2 x - (2D)_{px,py} (2D)_{py,px} x ; t = scratch full field
void orc_stagD2xx(void* lat, int nd, const double* g, double* r, const double* x, double m2, int parity, double* t) {
  Lat* l = (Lat*)lat;
  stag_sweep(l, nd, g, t, x, 0.0, x, 0.0, 1 - parity, 1.0);       // stagDP, rir := 0
  stag_sweep(l, nd, g, r, t, 0.0, x, 4.0 * m2, parity, -1.0);     // stagDM, rir := (4 m2) x
}
// eoReconstruct, stagD.nim:582-586: r.odd = (b.odd - D_oe r.even)/m
void orc_eoReconstruct(void* lat, int nd, const double* g, double* r, const double* b, double m) {
  Lat* l = (Lat*)lat;
  orc_stagD(lat, nd, g, r, r, 0.0, -1.0 / m, 0.0, 1);
#pragma omp parallel for
  for (int i = l->half * 6; i < l->vol * 6; i++) r[i] += b[i] / m;
}

// ---------------------------------------------------------------------------
// Field BLAS-1 on a subset (field/fieldET.nim:547-734)
// subset: 0 even, 1 odd, 2 all
// ---------------------------------------------------------------------------
static inline void sub_range(Lat* l, int sub, size_t& lo, size_t& hi) {
  lo = (sub == 1) ? (size_t)l->half * 6 : 0;
  hi = (sub == 0) ? (size_t)l->half * 6 : (size_t)l->vol * 6;
}
double orc_norm2(void* lat, const double* x, int sub) {
  size_t lo, hi; sub_range((Lat*)lat, sub, lo, hi);
  double s = 0;
#pragma omp parallel for reduction(+ : s)
  for (size_t i = lo; i < hi; i++) s += x[i] * x[i];
  return s;
}
double orc_redot(void* lat, const double* x, const double* y, int sub) {
  size_t lo, hi; sub_range((Lat*)lat, sub, lo, hi);
  double s = 0;
#pragma omp parallel for reduction(+ : s)
  for (size_t i = lo; i < hi; i++) s += x[i] * y[i];
  return s;
}

// ---------------------------------------------------------------------------
// CG  (src/solvers/cg.nim:55-272, cpNone), operator A = stagD2xx(m2) on `parity`
// x in/out (initial guess), b rhs.  Returns iterations; *r2out = final recursive r2.
// ---------------------------------------------------------------------------
int orc_cg_xx(void* lat, int nd, const double* g, double* x, const double* b, double m2, int parity,
              double r2req, int maxits, double* r2out, double* b2out) {
  Lat* l = (Lat*)lat;
  size_t n = (size_t)l->vol * 6, lo, hi; sub_range(l, parity, lo, hi);
  std::vector<double> r(n, 0.0), Ap(n, 0.0), p(n, 0.0), t(n, 0.0);
  double b2 = orc_norm2(lat, b, parity), r2 = 0.0;
  if (b2 == 0.0) {
    for (size_t i = lo; i < hi; i++) x[i] = 0;
    if (r2out) *r2out = 0; if (b2out) *b2out = 0;
    return 0;
  }
  orc_stagD2xx(lat, nd, g, Ap.data(), x, m2, parity, t.data());          // cg.nim:146
#pragma omp parallel for
  for (size_t i = lo; i < hi; i++) { r[i] = b[i] - Ap[i]; p[i] = 0; }   // :148-150
  r2 = orc_norm2(lat, r.data(), parity);
  double r2stop = r2req * b2;                                            // :155
  int itn = 0; double rzo = 1.0, rz;
  while (itn < maxits && r2 > r2stop) {                                  // :174
    rz = r2;                                                             // getRz, cpNone
    double beta = rz / rzo;                                              // :186
    rzo = rz;
    if (itn == 0) {
#pragma omp parallel for
      for (size_t i = lo; i < hi; i++) p[i] = r[i];                      // :190
    } else {
#pragma omp parallel for
      for (size_t i = lo; i < hi; i++) p[i] = r[i] + beta * p[i];        // :192
    }
    itn++;
    orc_stagD2xx(lat, nd, g, Ap.data(), p.data(), m2, parity, t.data()); // :200
    double pAp = orc_redot(lat, p.data(), Ap.data(), parity);            // :206
    double alpha = rz / pAp;                                             // :208
#pragma omp parallel for
    for (size_t i = lo; i < hi; i++) { x[i] += alpha * p[i]; r[i] -= alpha * Ap[i]; }  // :209-211
    r2 = orc_norm2(lat, r.data(), parity);                               // :213
  }
  if (r2out) *r2out = r2; if (b2out) *b2out = b2;
  return itn;
}

// solveXX with sbQex backend, stagSolve.nim:57-99: r := 0 then CG on A x = b
int orc_solveXX(void* lat, int nd, const double* g, double* r, const double* x, double m, int parity,
                double r2req, int maxits, double* r2out) {
  Lat* l = (Lat*)lat;
  memset(r, 0, sizeof(double) * (size_t)l->vol * 6);
  return orc_cg_xx(lat, nd, g, r, x, m * m, parity, r2req, maxits, r2out, nullptr);
}

struct OrcSolveStats { int iterations; int calls; double r2; double b2; };

// solveReconR, stagSolve.nim:141-176
static void solveReconR(Lat* l, int nd, const double* g, double* x, const double* b, double m,
                        double& r2req, int maxits, int& its, double b2e, double b2o) {
  size_t n = (size_t)l->vol * 6;
  double b2 = b2e + b2o, r2stop = r2req * b2, r2stop2 = 0.5 * r2stop;
  double r2stope = (b2o <= r2stop2) ? r2stop - b2o : r2stop2;
  if (b2e > r2stope) {
    std::vector<double> y(n, 0.0);
    r2req = r2stope / b2e;
    its += orc_solveXX(l, nd, g, y.data(), b, m, 0, r2req, maxits, nullptr);
    for (size_t i = 0; i < (size_t)l->half * 6; i++) y[i] *= 4;
    orc_Ddag(l, nd, g, x, y.data(), m);
    return;
  }
  double r2stopo = (b2e <= r2stop2) ? r2stop - b2e : r2stop2;
  if (b2o > r2stopo) {
    std::vector<double> y(n, 0.0);
    r2req = r2stopo / b2o;
    its += orc_solveXX(l, nd, g, y.data(), b, m, 1, r2req, maxits, nullptr);
    for (size_t i = (size_t)l->half * 6; i < n; i++) y[i] *= 4;
    orc_Ddag(l, nd, g, x, y.data(), m);
    return;
  }
}
// solveReconL, stagSolve.nim:179-208
static void solveReconL(Lat* l, int nd, const double* g, double* x, const double* b, double m,
                        double& r2req, int maxits, int& its, double b2e, double b2o) {
  size_t n = (size_t)l->vol * 6;
  std::vector<double> d(n, 0.0);
  orc_Ddag(l, nd, g, d.data(), b, m);
  memset(x, 0, sizeof(double) * n);
  double d2e = orc_norm2(l, d.data(), 0);
  r2req = 0.99 * r2req * (b2e + b2o) * m * m / d2e;
  its += orc_solveXX(l, nd, g, x, d.data(), m, 0, r2req, maxits, nullptr);
  for (size_t i = 0; i < (size_t)l->half * 6; i++) x[i] *= 4;
  orc_eoReconstruct(l, nd, g, x, b, m);
}
// solveInner, stagSolve.nim:211-222
static void solveInner(Lat* l, int nd, const double* g, double* x, const double* b, double m,
                       double& r2req, int maxits, int& its, double b2e, double b2o) {
  double b2 = b2e + b2o, r2stop = r2req * b2, r2stop2 = 0.5 * r2stop;
  double r2stope = (b2o <= r2stop2) ? r2stop - b2o : r2stop2;
  double r2stopo = (b2e <= r2stop2) ? r2stop - b2e : r2stop2;
  if (b2e <= r2stope || b2o <= r2stopo || m == 0.0) solveReconR(l, nd, g, x, b, m, r2req, maxits, its, b2e, b2o);
  else solveReconL(l, nd, g, x, b, m, r2req, maxits, its, b2e, b2o);
}
// solve, stagSolve.nim:224-294.  Full-lattice D x = b.
void orc_solve(void* lat, int nd, const double* g, double* x, const double* b, double m,
               double r2req0, int maxits0, int usePrevSoln, OrcSolveStats* st) {
  Lat* l = (Lat*)lat;
  size_t n = (size_t)l->vol * 6;
  double b2 = orc_norm2(lat, b, 2);
  double r2stop = r2req0 * b2;
  std::vector<double> r(n), y(n, 0.0);
  if (usePrevSoln) {
    orc_D(lat, nd, g, r.data(), x, m);
    for (size_t i = 0; i < n; i++) r[i] = b[i] - r[i];
  } else {
    memset(x, 0, sizeof(double) * n);
    memcpy(r.data(), b, sizeof(double) * n);
  }
  double r2e = orc_norm2(lat, r.data(), 0), r2o = orc_norm2(lat, r.data(), 1);
  double r2 = r2e + r2o;
  int its = 0, calls = 0;
  while (r2 > r2stop) {
    int maxits = maxits0 - its;
    if (maxits <= 0) break;
    double r2req = r2stop / r2;
    solveInner(l, nd, g, y.data(), r.data(), m, r2req, maxits, its, r2e, r2o);
    calls++;
    for (size_t i = 0; i < n; i++) x[i] += y[i];
    orc_D(lat, nd, g, r.data(), x, m);
    for (size_t i = 0; i < n; i++) r[i] = b[i] - r[i];
    r2e = orc_norm2(lat, r.data(), 0); r2o = orc_norm2(lat, r.data(), 1);
    r2 = r2e + r2o;
  }
  if (st) { st->iterations = its; st->calls = calls; st->r2 = (b2 > 0 ? r2 / b2 : 0.0); st->b2 = b2; }
}

// ---------------------------------------------------------------------------
// Multi-shift CG  (src/solvers/cgm.nim:84-316, cpNone: z=r, q=ps[0], LAp=Ap), new solution
// (b2 < 0 branch, :166-172).  Solves (A + sigma_k) xs[k] = b on `parity`, A = stagD2xx(m2),
// sg[0] = 0, sg[k>0] = sigmas[k] (:107-110).  xs: nmass full-lattice vectors, contiguous.
// Returns iterations; *r2out = final r2 of the unshifted system, *b2out = |b|^2 on the subset.
// ---------------------------------------------------------------------------
int orc_cgm_xx(void* lat, int nd, const double* g, double* xs, const double* b, double m2,
               const double* sigmas, int nmass, int parity, double r2req, int maxits,
               double* r2out, double* b2out) {
  Lat* l = (Lat*)lat;
  const size_t n = (size_t)l->vol * 6;
  size_t lo, hi; sub_range(l, parity, lo, hi);
  std::vector<double> r(n, 0.0), Ap(n, 0.0), t(n, 0.0), ps((size_t)nmass * n, 0.0);
  std::vector<double> sg(nmass), zi(nmass, 1.0), zim1(nmass, 1.0);
  for (int m = 0; m < nmass; m++) sg[m] = (m == 0) ? 0.0 : sigmas[m];
#pragma omp parallel for
  for (size_t i = lo; i < hi; i++) r[i] = b[i];                              // :168
  for (int m = 0; m < nmass; m++)
    for (size_t i = lo; i < hi; i++) xs[(size_t)m * n + i] = 0.0;            // :169
  const double b2 = orc_norm2(lat, b, parity);
  double r2 = b2;
  const double r2stop = r2req * b2;                                          // :184
  int itn = 0;
  if (r2 > r2stop) {
    double alpha = 0, beta = 0, alphaim1 = -1.0, betaim1 = 0.0, r2i = r2, r2ip1 = 0, r2r = 0;   // :199-211
    double* p = ps.data();
    for (int m = 0; m < nmass; m++)
      for (size_t i = lo; i < hi; i++) ps[(size_t)m * n + i] = r[i];         // q := z ; ps[m] := p  (:205-218)
    bool continuing = true;
    while (continuing) {                                                     // :223
      for (int m = 0; m < nmass; m++) {
        if (m == 0) {
          orc_stagD2xx(lat, nd, g, Ap.data(), p, m2, parity, t.data());      // :228
          itn++;
          const double qLAp = orc_redot(lat, p, Ap.data(), parity);          // :233
          alpha = (qLAp != 0.0) ? r2i / qLAp : 0.0;
          double* x = xs;
#pragma omp parallel for
          for (size_t i = lo; i < hi; i++) { r[i] -= alpha * Ap[i]; x[i] += alpha * p[i]; }   // :237-239
          r2ip1 = orc_norm2(lat, r.data(), parity);                          // :245-248
          r2r = r2ip1;
          beta = (r2i != 0.0) ? r2ip1 / r2i : 0.0;                           // :252-254
          continuing = (itn < maxits) && (r2r > r2stop);                     // :256
          if (continuing) {
#pragma omp parallel for
            for (size_t i = lo; i < hi; i++) p[i] = r[i] + beta * p[i];      // :258
          }
        } else {                                                             // :260-274
          double zip1d = alpha * betaim1 * (zim1[m] - zi[m]);
          zip1d += zim1[m] * alphaim1 * (1.0 + sg[m] * alpha);
          const double zip1 = (zip1d != 0.0) ? zi[m] * zim1[m] * alphaim1 / zip1d : 0.0;
          const double zr = (zi[m] != 0.0) ? zip1 / zi[m] : 0.0;
          double* xm = xs + (size_t)m * n;
          double* pm = ps.data() + (size_t)m * n;
          const double cx = alpha * zr;
#pragma omp parallel for
          for (size_t i = lo; i < hi; i++) xm[i] += cx * pm[i];
          if (continuing) {
            const double cp = beta * zr * zr;
#pragma omp parallel for
            for (size_t i = lo; i < hi; i++) pm[i] = zip1 * r[i] + cp * pm[i];
            zim1[m] = zi[m]; zi[m] = zip1;
          }
        }
      }
      alphaim1 = alpha; betaim1 = beta; r2i = r2ip1;                          // :276
      r2 = r2r;
    }
  }
  if (r2out) *r2out = r2;
  if (b2out) *b2out = b2;
  return itn;
}

// solve(s, xs, b, ms, sp0), stagSolve.nim:347-446: xs[k] = Ddag(m_k) (D Ddag)^-1 b through the
// even (or odd) multi-shift solve; shifts[0] = ms[0], shifts[k] = 4 (ms[k]^2 - ms[0]^2).
// Returns -1 if neither parity qualifies for a solve (the reference would raise on subset "").
int orc_solve_multi(void* lat, int nd, const double* g, double* xs, const double* b, const double* ms,
                    int nmass, double r2req0, int maxits0, OrcSolveStats* st) {
  Lat* l = (Lat*)lat;
  const size_t n = (size_t)l->vol * 6, nh = (size_t)l->half * 6;
  const double mass = ms[0];
  double b2 = orc_norm2(lat, b, 2), b2e = orc_norm2(lat, b, 0), b2o = orc_norm2(lat, b, 1);
  std::vector<double> r(b, b + n), ys((size_t)nmass * n, 0.0), xt(n, 0.0), shifts(nmass);
  double r2 = b2e + b2o;
  const double r2stop = r2req0 * b2;
  for (int m = 0; m < nmass; m++) shifts[m] = (m == 0) ? ms[m] : 4.0 * (ms[m] * ms[m] - mass * mass);
  memset(xs, 0, sizeof(double) * n * nmass);
  int its = 0, calls = 0, subset = -1;
  while (r2 > r2stop) {
    const int maxits = maxits0 - its;
    if (maxits <= 0) break;
    double r2req = r2stop;
    const double r2stop2 = 0.5 * r2req;
    const double r2stope = (b2o <= r2stop2) ? r2req - b2o : r2stop2;
    const double r2stopo = (b2e <= r2stop2) ? r2req - b2e : r2stop2;
    if (b2e > r2stope) { r2req = r2stope / b2e; subset = 0; }
    else if (b2o > r2stopo) { r2req = r2stopo / b2o; subset = 1; }
    if (subset < 0) return -1;
    std::fill(ys.begin(), ys.end(), 0.0);
    its += orc_cgm_xx(lat, nd, g, ys.data(), r.data(), mass * mass, shifts.data(), nmass, subset, r2req, maxits,
                      nullptr, nullptr);
    calls++;
    const size_t lo = subset == 0 ? 0 : nh, hi = subset == 0 ? nh : n;
    for (int m = 0; m < nmass; m++)
      for (size_t i = lo; i < hi; i++) xs[(size_t)m * n + i] += 4.0 * ys[(size_t)m * n + i];
    orc_Ddag(lat, nd, g, xt.data(), xs, mass);
    orc_D(lat, nd, g, r.data(), xt.data(), mass);
    for (size_t i = 0; i < n; i++) r[i] = b[i] - r[i];
    b2e = orc_norm2(lat, r.data(), 0); b2o = orc_norm2(lat, r.data(), 1);
    r2 = b2e + b2o;
  }
  for (int m = 0; m < nmass; m++) {                                          // :431-437
    if (m != 0 || calls == 0) orc_Ddag(lat, nd, g, xt.data(), xs + (size_t)m * n, ms[m]);
    memcpy(xs + (size_t)m * n, xt.data(), sizeof(double) * n);
  }
  if (st) { st->iterations = its; st->calls = calls; st->r2 = (b2 > 0 ? r2 / b2 : 0.0); st->b2 = b2; }
  return 0;
}

// ---------------------------------------------------------------------------
// Link smearing: fat7 + Lepage + Naik (src/gauge/fat7l.nim:24-162), HISQ
// (src/physics/hisqLinks.nim:9-43), nHYP (src/gauge/hypsmear.nim:49-162,
// src/gauge/smearutil.nim:3-22).  Matrix fields are [vol] of M3; a link set is
// 4 contiguous fields [mu][vol][18].  The restatement keeps the reference's
// structure: shifted copies first (Shifter ^*), then site-local products.
// ---------------------------------------------------------------------------
typedef std::vector<M3> MF;
static MF mf_load(Lat* l, const double* p) {
  MF f(l->vol);
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) f[i] = m3_load(p + (size_t)i * 18);
  return f;
}
static void mf_store(Lat* l, double* p, const MF& f) {
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) m3_store(p + (size_t)i * 18, f[i]);
}
// newShifter(f, mu, +1) ^* f : r(x) = f(x + mu);  len = -1 : r(x) = f(x - mu)
static MF mf_shift(Lat* l, const MF& f, int mu, int len) {
  MF r(l->vol);
  const std::vector<int>& t = len > 0 ? l->fwd[mu][0] : l->bwd[mu][0];
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) r[i] = f[t[i]];
  return r;
}
static inline void m3_axpy(M3& y, double a, const M3& x) {
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) y.e[i][j] += a * x.e[i][j];
}
static inline M3 m3_scale(double a, const M3& x) {
  M3 r; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.e[i][j] = a * x.e[i][j]; return r;
}

// computeGenStaple, fat7l.nim:24-76.  staple == nullptr: the `bool` overload (not saved).
static void gen_staple(Lat* l, MF* staple, int mu, int nu, const MF& link, bool link_is_gauge_mu, double coef,
                       const MF* gauge, MF& flmu) {
  (void)link_is_gauge_mu;
  const MF ts0 = mf_shift(l, link, nu, +1);        // link(x+nu)
  const MF ts1 = mf_shift(l, gauge[nu], mu, +1);   // gauge[nu](x+mu)
  MF tmat2(l->vol);
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) {
    const M3 up = m3_mul(m3_mul(gauge[nu][i], ts0[i]), m3_adj(ts1[i]));       // :43,48
    if (staple) (*staple)[i] = up; else m3_axpy(flmu[i], coef, up);
    tmat2[i] = m3_mul(m3_mul(m3_adj(gauge[nu][i]), link[i]), ts1[i]);         // :54
  }
  const MF ts2 = mf_shift(l, tmat2, nu, -1);                                  // :55
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) {
    if (staple) { m3_axpy((*staple)[i], 1.0, ts2[i]); m3_axpy(flmu[i], coef, (*staple)[i]); }   // :60-61
    else m3_axpy(flmu[i], coef, ts2[i]);                                      // :64
  }
}

// makeImpLinks, fat7l.nim:78-162.  coef = {oneLink, threeStaple, fiveStaple, sevenStaple, lepage}.
// gfl / ll may be NULL when naik == 0.
void orc_makeImpLinks(void* lat, const double* gf_, const double* coef, double* fl_, const double* gfl_, double naik,
                      double* ll_) {
  Lat* l = (Lat*)lat;
  const size_t fs = (size_t)l->vol * 18;
  MF gf[4], fl[4];
  for (int d = 0; d < 4; d++) gf[d] = mf_load(l, gf_ + d * fs);
  double coef1 = coef[0];
  const double coef3 = coef[1], coef5 = coef[2], coef7 = coef[3], coefL = coef[4];
  const bool have5 = coef5 != 0.0 || coef7 != 0.0 || coefL != 0.0;
  const bool have3 = coef3 != 0.0 || have5;
  coef1 -= 6.0 * coefL;                                                       // :108
  for (int dir = 0; dir < 4; dir++) {
    fl[dir].resize(l->vol);
#pragma omp parallel for
    for (int i = 0; i < l->vol; i++) fl[dir][i] = m3_scale(coef1, gf[dir][i]);  // :135
    if (!have3) continue;
    for (int nu = 0; nu < 4; nu++) {
      if (nu == dir) continue;
      MF staple(l->vol), tempmat1(l->vol);
      gen_staple(l, &staple, dir, nu, gf[dir], true, coef3, gf, fl[dir]);     // :139
      if (coefL != 0.0) gen_staple(l, nullptr, dir, nu, staple, false, coefL, gf, fl[dir]);  // :141-143
      if (coef5 != 0.0 || coef7 != 0.0) {
        for (int rho = 0; rho < 4; rho++) {
          if (rho == dir || rho == nu) continue;
          gen_staple(l, &tempmat1, dir, rho, staple, false, coef5, gf, fl[dir]);  // :147
          if (coef7 != 0.0) {
            for (int sig = 0; sig < 4; sig++) {
              if (sig == dir || sig == nu || sig == rho) continue;
              gen_staple(l, nullptr, dir, sig, tempmat1, false, coef7, gf, fl[dir]);  // :152
            }
          }
        }
      }
    }
  }
  if (naik != 0.0) {                                                          // :156-167
    for (int dir = 0; dir < 4; dir++) {
      const MF w = mf_load(l, gfl_ + dir * fs);
      const MF s1 = mf_shift(l, w, dir, +1);
      MF t(l->vol);
#pragma omp parallel for
      for (int i = 0; i < l->vol; i++) t[i] = m3_mul(w[i], s1[i]);
      const MF s2 = mf_shift(l, t, dir, +1);
      MF r(l->vol);
#pragma omp parallel for
      for (int i = 0; i < l->vol; i++) r[i] = m3_scale(naik, m3_mul(w[i], s2[i]));
      mf_store(l, ll_ + dir * fs, r);
    }
  }
  for (int d = 0; d < 4; d++) mf_store(l, fl_ + d * fs, fl[d]);
}

// projectU of every site of n fields (hisqLinks.nim:35-37, hypsmear.nim:28-34)
void orc_projectU_field(void* lat, const double* in, double* out, int nfields) {
  Lat* l = (Lat*)lat;
  const size_t n = (size_t)l->vol * nfields;
#pragma omp parallel for
  for (size_t i = 0; i < n; i++) m3_store(out + i * 18, projectU(m3_load(in + i * 18)));
}

// HisqCoefs.smear, hisqLinks.nim:32-39.  c = {fat7first[5], fat7second[5], naik}.
void orc_hisq_smear(void* lat, const double* g, const double* c, double* fl, double* ll) {
  Lat* l = (Lat*)lat;
  const size_t n = (size_t)l->vol * 18 * 4;
  std::vector<double> t1(n), t2(n);
  orc_makeImpLinks(lat, g, c, t1.data(), nullptr, 0.0, nullptr);
  orc_projectU_field(lat, t1.data(), t2.data(), 4);
  orc_makeImpLinks(lat, t2.data(), c + 5, fl, t2.data(), c[10], ll);
}

// symStaple, smearutil.nim:3-22: s += alp*[ g1 * s2 * s1^+  +  (g1^+ g2 s1)(x - d1) ],
// s1 = g1 shifted from +d2, s2 = g2 shifted from +d1 (d1 / d2 = direction of g1 / g2).
static void sym_staple(Lat* l, MF& s, double alp, const MF& g1, int d1, const MF& g2, int d2) {
  const MF s1 = mf_shift(l, g1, d2, +1), s2 = mf_shift(l, g2, d1, +1);
  MF tm(l->vol);
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) tm[i] = m3_mul(m3_mul(m3_adj(g1[i]), g2[i]), s1[i]);
  const MF sm = mf_shift(l, tm, d1, -1);
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) {
    m3_axpy(s[i], alp, m3_mul(m3_mul(g1[i], s2[i]), m3_adj(s1[i])));
    m3_axpy(s[i], alp, sm[i]);
  }
}
static void mf_project(MF& r, const MF& x) {
#pragma omp parallel for
  for (size_t i = 0; i < x.size(); i++) r[i] = projectU(x[i]);
}

// HypCoefs.smear = forward part of smearGetForce, hypsmear.nim:49-162.  alpha = {alpha1, alpha2, alpha3}.
// Optional outputs (may be NULL): l1o/l2o [4][4][vol][18] projected level-1/2 links (diagonal unused).
void orc_hyp_smear(void* lat, const double* g, const double* alpha, double* fl_) {
  Lat* l = (Lat*)lat;
  const size_t fs = (size_t)l->vol * 18;
  MF gf[4], l1[4][4], l2[4][4];
  for (int d = 0; d < 4; d++) gf[d] = mf_load(l, g + d * fs);
  const double alp1 = alpha[0] / 2.0, alp2 = alpha[1] / 4.0, alp3 = alpha[2] / 6.0;
  const double ma1 = 1 - alpha[0], ma2 = 1 - alpha[1], ma3 = 1 - alpha[2];
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {
    if (nu == mu) continue;
    MF x(l->vol);
#pragma omp parallel for
    for (int i = 0; i < l->vol; i++) x[i] = m3_scale(ma1, gf[mu][i]);        // :107
    sym_staple(l, x, alp1, gf[nu], nu, gf[mu], mu);                           // :108-109
    l1[mu][nu].resize(l->vol);
    mf_project(l1[mu][nu], x);
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {
    if (nu == mu) continue;
    MF x(l->vol);
#pragma omp parallel for
    for (int i = 0; i < l->vol; i++) x[i] = m3_scale(ma2, gf[mu][i]);        // :117
    for (int a = 0; a < 4; a++) {
      if (a == mu || a == nu) continue;
      const int b = 1 + 2 + 3 - mu - nu - a;
      sym_staple(l, x, alp2, l1[a][b], a, l1[mu][b], mu);                     // :129-130
    }
    l2[mu][nu].resize(l->vol);
    mf_project(l2[mu][nu], x);
  }
  for (int mu = 0; mu < 4; mu++) {
    MF x(l->vol);
#pragma omp parallel for
    for (int i = 0; i < l->vol; i++) x[i] = m3_scale(ma3, gf[mu][i]);        // :136
    for (int nu = 0; nu < 4; nu++) {
      if (nu == mu) continue;
      sym_staple(l, x, alp3, l2[nu][mu], nu, l2[mu][nu], mu);                 // :147-148
    }
    MF r(l->vol);
    mf_project(r, x);
    mf_store(l, fl_ + mu * fs, r);
  }
}

// ---------------------------------------------------------------------------
// nHYP smearing force: smearGetForce's smearedForce (src/gauge/hypsmear.nim:163-262),
// symStapleDeriv (src/gauge/smearutil.nim:24-54), projectUderiv (src/maths/matrixFunctions.nim:321-357),
// inverse (src/maths/matinv.nim:89-113), adjugate + sylsolve (src/maths/projUderiv.nim:8-38,96-148).
// ---------------------------------------------------------------------------
static inline M3 m3_add(const M3& a, const M3& b) { M3 r = a; m3_axpy(r, 1.0, b); return r; }
static inline M3 m3_inverse(const M3& x) {   // matinv.nim:89-113
  const cplx x00 = x.e[0][0], x01 = x.e[0][1], x02 = x.e[0][2], x10 = x.e[1][0], x11 = x.e[1][1], x12 = x.e[1][2],
             x20 = x.e[2][0], x21 = x.e[2][1], x22 = x.e[2][2];
  const cplx det0 = x00 * x11 - x01 * x10, det1 = x02 * x10 - x00 * x12, det2 = x01 * x12 - x02 * x11;
  const cplx det = det0 * x22 + det1 * x21 + det2 * x20;
  const cplx idet = 1.0 / det;
  M3 r;
  r.e[0][0] = idet * (x11 * x22 - x12 * x21); r.e[0][1] = idet * (x21 * x02 - x22 * x01); r.e[0][2] = idet * det2;
  r.e[1][0] = idet * (x12 * x20 - x10 * x22); r.e[1][1] = idet * (x22 * x00 - x20 * x02); r.e[1][2] = idet * det1;
  r.e[2][0] = idet * (x10 * x21 - x11 * x20); r.e[2][1] = idet * (x20 * x01 - x21 * x00); r.e[2][2] = idet * det0;
  return r;
}
static inline M3 m3_adjugate(const M3& x) {  // projUderiv.nim:17-35
  M3 r;
  r.e[0][0] = x.e[1][1] * x.e[2][2] - x.e[1][2] * x.e[2][1]; r.e[0][1] = x.e[2][1] * x.e[0][2] - x.e[2][2] * x.e[0][1];
  r.e[0][2] = x.e[0][1] * x.e[1][2] - x.e[0][2] * x.e[1][1]; r.e[1][0] = x.e[1][2] * x.e[2][0] - x.e[1][0] * x.e[2][2];
  r.e[1][1] = x.e[2][2] * x.e[0][0] - x.e[2][0] * x.e[0][2]; r.e[1][2] = x.e[0][2] * x.e[1][0] - x.e[0][0] * x.e[1][2];
  r.e[2][0] = x.e[1][0] * x.e[2][1] - x.e[1][1] * x.e[2][0]; r.e[2][1] = x.e[2][0] * x.e[0][1] - x.e[2][1] * x.e[0][0];
  r.e[2][2] = x.e[0][0] * x.e[1][1] - x.e[0][1] * x.e[1][0];
  return r;
}
// solves A X + X A = C, projUderiv.nim:122-146
static inline M3 m3_sylsolve(const M3& a, const M3& c) {
  const M3 ad = m3_adjugate(a);
  const cplx t = a.e[0][0] + a.e[1][1] + a.e[2][2];
  const cplx s = ad.e[0][0] + ad.e[1][1] + ad.e[2][2];
  const cplx r = a.e[0][0] * ad.e[0][0] + a.e[0][1] * ad.e[1][0] + a.e[0][2] * ad.e[2][0];
  const M3 ac = m3_mul(a, c), ca = m3_mul(c, a), aca = m3_mul(ac, a), adc = m3_mul(ad, c), cad = m3_mul(c, ad),
           adcad = m3_mul(adc, ad);
  const cplx c2 = 1.0 / (2.0 * (s * t - r)), c0 = c2 * (s + t * t), c1 = c2 * (t / r), c4 = c2 * t;
  M3 x;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++)
    x.e[i][j] = c0 * c.e[i][j] + c1 * adcad.e[i][j] + c2 * (aca.e[i][j] - adc.e[i][j] - cad.e[i][j]) -
                c4 * (ac.e[i][j] + ca.e[i][j]);
  return x;
}
// projectUderiv(r, u, x, chain), matrixFunctions.nim:329-351
static inline M3 projectUderiv(const M3& u, const M3& x, const M3& chain) {
  M3 t = m3_mul(m3_adj(x), x);
  for (int i = 0; i < 3; i++) t.e[i][i] += 1e-20;
  const M3 z = rsqrtPH3(t);
  const M3 y = m3_inverse(z);
  M3 r = m3_mul(chain, z);
  const M3 t1 = m3_mul(m3_adj(u), r);
  const M3 t2 = m3_sylsolve(y, t1);
  const M3 t3 = m3_add(t2, m3_adj(t2));
  m3_axpy(r, -1.0, m3_mul(x, t3));
  return r;
}
void orc_projectUderiv(const double* x, const double* chain, double* out) {   // :353-357
  const M3 X = m3_load(x);
  m3_store(out, projectUderiv(projectU(X), X, m3_load(chain)));
}
// symStapleDeriv, smearutil.nim:24-54.  d1 / d2 = direction of g1 / g2; c = chain (direction d2).
static void sym_staple_deriv(Lat* l, MF& f1, MF& f2, const MF& g1, int d1, const MF& g2, int d2, const MF& c) {
  const MF s1 = mf_shift(l, g1, d2, +1), s2 = mf_shift(l, g2, d1, +1), s = mf_shift(l, c, d1, +1);
  MF tm1(l->vol), tm2(l->vol);
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) {
    tm1[i] = m3_mul(m3_mul(m3_adj(g1[i]), c[i]), s1[i]);
    tm2[i] = m3_add(m3_mul(m3_mul(m3_adj(g2[i]), g1[i]), s[i]), m3_mul(m3_mul(m3_adj(c[i]), g1[i]), s2[i]));
  }
  const MF sm1 = mf_shift(l, tm1, d1, -1), sm2 = mf_shift(l, tm2, d2, -1);
#pragma omp parallel for
  for (int i = 0; i < l->vol; i++) {
    m3_axpy(f1[i], 1.0, m3_mul(m3_mul(g2[i], s1[i]), m3_adj(s[i])));
    m3_axpy(f1[i], 1.0, m3_mul(m3_mul(c[i], s1[i]), m3_adj(s2[i])));
    m3_axpy(f2[i], 1.0, m3_mul(m3_mul(g1[i], s[i]), m3_adj(s1[i])));
    m3_axpy(f2[i], 1.0, sm1[i]);
    m3_axpy(f1[i], 1.0, sm2[i]);
  }
}
// test hook: f1, f2 += symStapleDeriv(g1 (dir d1), g2 (dir d2), chain c)
void orc_sym_staple_deriv(void* lat, double* f1_, double* f2_, const double* g1_, int d1, const double* g2_, int d2,
                          const double* c_) {
  Lat* l = (Lat*)lat;
  MF f1 = mf_load(l, f1_), f2 = mf_load(l, f2_);
  sym_staple_deriv(l, f1, f2, mf_load(l, g1_), d1, mf_load(l, g2_), d2, mf_load(l, c_));
  mf_store(l, f1_, f1); mf_store(l, f2_, f2);
}
static void mf_scale(MF& f, double a) {
#pragma omp parallel for
  for (size_t i = 0; i < f.size(); i++) f[i] = m3_scale(a, f[i]);
}
static void mf_axpy(MF& y, double a, const MF& x) {
#pragma omp parallel for
  for (size_t i = 0; i < y.size(); i++) m3_axpy(y[i], a, x[i]);
}

// smear (forward, keepProj branch) + smearedForce(f, chain), hypsmear.nim:49-262.  fl_ (may be NULL) gets the
// smeared links; f_ the force with respect to the thin links.
void orc_hyp_force(void* lat, const double* g, const double* alpha, const double* chain_, double* f_, double* fl_) {
  Lat* l = (Lat*)lat;
  const size_t fs = (size_t)l->vol * 18;
  const int V = l->vol;
  MF gf[4], l1[4][4], l2[4][4], l1x[4][4], l2x[4][4], flx[4];
  for (int d = 0; d < 4; d++) gf[d] = mf_load(l, g + d * fs);
  const double alp1 = alpha[0] / 2.0, alp2 = alpha[1] / 4.0, alp3 = alpha[2] / 6.0;
  const double ma1 = 1 - alpha[0], ma2 = 1 - alpha[1], ma3 = 1 - alpha[2];
  auto scaled = [&](double a, const MF& x) { MF r(V); for (int i = 0; i < V; i++) r[i] = m3_scale(a, x[i]); return r; };
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {
    if (nu == mu) continue;
    l1x[mu][nu] = scaled(ma1, gf[mu]);
    sym_staple(l, l1x[mu][nu], alp1, gf[nu], nu, gf[mu], mu);
    l1[mu][nu].resize(V); mf_project(l1[mu][nu], l1x[mu][nu]);
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {
    if (nu == mu) continue;
    l2x[mu][nu] = scaled(ma2, gf[mu]);
    for (int a = 0; a < 4; a++) {
      if (a == mu || a == nu) continue;
      const int b = 6 - mu - nu - a;
      sym_staple(l, l2x[mu][nu], alp2, l1[a][b], a, l1[mu][b], mu);
    }
    l2[mu][nu].resize(V); mf_project(l2[mu][nu], l2x[mu][nu]);
  }
  for (int mu = 0; mu < 4; mu++) {
    flx[mu] = scaled(ma3, gf[mu]);
    for (int nu = 0; nu < 4; nu++) if (nu != mu) sym_staple(l, flx[mu], alp3, l2[nu][mu], nu, l2[mu][nu], mu);
    if (fl_) { MF r(V); mf_project(r, flx[mu]); mf_store(l, fl_ + mu * fs, r); }
  }
  // ---- smearedForce, :163-262 ----
  MF f[4], fc[4], fl1[4][4], fl2[4][4];
  const M3 zero = m3_scale(0.0, gf[0][0]);
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) if (nu != mu) { fl1[mu][nu].assign(V, zero); fl2[mu][nu].assign(V, zero); }
  for (int mu = 0; mu < 4; mu++) {                                             // :185-186
    const MF ch = mf_load(l, chain_ + mu * fs);
    fc[mu].resize(V);
#pragma omp parallel for
    for (int i = 0; i < V; i++) fc[mu][i] = projectUderiv(projectU(flx[mu][i]), flx[mu][i], ch[i]);
  }
  for (int mu = 0; mu < 4; mu++) { f[mu] = scaled(ma3, fc[mu]); mf_scale(fc[mu], alp3); }   // :188-190
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {                          // :191-205
    if (nu == mu) continue;
    sym_staple_deriv(l, fl2[nu][mu], fl2[mu][nu], l2[nu][mu], nu, l2[mu][nu], mu, fc[mu]);
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {                          // :209-215
    if (nu == mu) continue;
    MF& x = fl2[mu][nu];
#pragma omp parallel for
    for (int i = 0; i < V; i++) x[i] = projectUderiv(l2[mu][nu][i], l2x[mu][nu][i], x[i]);
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {                          // :217-221
    if (nu == mu) continue;
    mf_axpy(f[mu], ma2, fl2[mu][nu]);
    mf_scale(fl2[mu][nu], alp2);
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {                          // :222-240
    if (nu == mu) continue;
    for (int a = 0; a < 4; a++) {
      if (a == mu || a == nu) continue;
      const int b = 6 - mu - nu - a;
      sym_staple_deriv(l, fl1[a][b], fl1[mu][b], l1[a][b], a, l1[mu][b], mu, fl2[mu][nu]);
    }
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {                          // :244-251
    if (nu == mu) continue;
    MF& x = fl1[mu][nu];
#pragma omp parallel for
    for (int i = 0; i < V; i++) x[i] = projectUderiv(l1[mu][nu][i], l1x[mu][nu][i], x[i]);
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {                          // :253-257
    if (nu == mu) continue;
    mf_axpy(f[mu], ma1, fl1[mu][nu]);
    mf_scale(fl1[mu][nu], alp1);
  }
  for (int mu = 0; mu < 4; mu++) for (int nu = 0; nu < 4; nu++) {                          // :258-264
    if (nu == mu) continue;
    sym_staple_deriv(l, f[nu], f[mu], gf[nu], nu, gf[mu], mu, fl1[mu][nu]);
  }
  for (int mu = 0; mu < 4; mu++) mf_store(l, f_ + mu * fs, f[mu]);
}

// ---------------------------------------------------------------------------
// fat7 + Lepage + Naik derivative (src/gauge/fat7lderiv.nim:122-335) and the HISQ smearing force
// (src/gauge/hisqsmear.nim:51-84).  Built, like the reference, from symStaple / symStapleDeriv.
// deriv / llderiv ACCUMULATE (the reference's `+=`).
// ---------------------------------------------------------------------------
static void fat7l_deriv(Lat* l, MF* deriv, const MF* gauge, const MF* mid, const double* coef, MF* llderiv,
                        const MF* llgauge, const MF* llmid, double naik) {
  const int V = l->vol;
  const double coefL = coef[4], coef1 = coef[0] - 6 * coefL, coef3 = coef[1], coef5 = coef[2], coef7 = coef[3];
  const bool have5 = coef5 != 0.0 || coef7 != 0.0 || coefL != 0.0;
  const bool have3 = coef3 != 0.0 || have5;
  const M3 zero = m3_scale(0.0, gauge[0][0]);
  MF stpl3[4], mid5[4];
  for (int sig = 0; sig < 4; sig++) {
    if (have5) {
      for (int mu = 0; mu < 4; mu++) {
        if (mu == sig) continue;
        stpl3[mu].assign(V, zero);
        sym_staple(l, stpl3[mu], 1.0, gauge[mu], mu, gauge[sig], sig);      // :185-186
        mid5[mu].assign(V, zero);
      }
      for (int rho = 0; rho < 4; rho++) {
        if (rho == sig) continue;
        MF stpl5(V, zero), mid3(V, zero);
        if (coef7 != 0.0) {
          for (int mu = 0; mu < 4; mu++) {
            if (mu == sig || mu == rho) continue;
            for (int nu = 0; nu < 4; nu++) {
              if (nu == sig || nu == rho || nu == mu) continue;
              sym_staple(l, stpl5, 1.0, gauge[nu], nu, stpl3[mu], sig);     // :202-203
            }
          }
          mf_scale(stpl5, coef7);
        }
        if (coefL != 0.0) mf_axpy(stpl5, coefL, stpl3[rho]);                // :212
        if (coefL != 0.0 || coef7 != 0.0)
          sym_staple_deriv(l, deriv[rho], mid3, gauge[rho], rho, stpl5, sig, mid[sig]);   // :218-220
        if (coefL != 0.0) mf_axpy(mid5[rho], coefL, mid3);                  // :224
        mf_scale(mid3, coef7);                                              // :228
        mf_axpy(mid3, coef5, mid[sig]);                                     // :230
        for (int mu = 0; mu < 4; mu++) {
          if (mu == sig || mu == rho) continue;
          for (int nu = 0; nu < 4; nu++) {
            if (nu == sig || nu == rho || nu == mu) continue;
            sym_staple_deriv(l, deriv[mu], mid5[nu], gauge[mu], mu, stpl3[nu], sig, mid3);   // :239-241
          }
        }
      }
    }
    if (have3) {
      MF mid3(V, zero);
      for (int mu = 0; mu < 4; mu++) {
        if (mu == sig) continue;
        MF stpl5(V);
        for (int i = 0; i < V; i++) stpl5[i] = m3_scale(coef3, mid[sig][i]);
        if (have5) mf_axpy(stpl5, 1.0, mid5[mu]);                           // :250-256
        sym_staple_deriv(l, deriv[mu], mid3, gauge[mu], mu, gauge[sig], sig, stpl5);   // :259-261
      }
      mf_axpy(deriv[sig], 1.0, mid3);                                       // :265
    }
    if (coef1 != 0.0) mf_axpy(deriv[sig], coef1, mid[sig]);                 // :269
  }
  if (naik != 0.0) {                                                        // :300-326
    for (int mu = 0; mu < 4; mu++) {
      const MF& U = llgauge[mu];
      const MF& M = llmid[mu];
      const MF s1a = mf_shift(l, U, mu, +1);
      MF tm1(V), tm2(V);
      for (int i = 0; i < V; i++) { tm1[i] = m3_mul(m3_adj(U[i]), M[i]); tm2[i] = m3_mul(m3_adj(s1a[i]), m3_adj(U[i])); }
      const MF sm1 = mf_shift(l, tm1, mu, -1);
      const MF s1b = mf_shift(l, tm2, mu, +1);
      for (int i = 0; i < V; i++) {
        m3_axpy(llderiv[mu][i], naik, m3_mul(sm1[i], m3_adj(s1a[i])));
        tm1[i] = m3_mul(m3_adj(U[i]), sm1[i]);
      }
      const MF sm2 = mf_shift(l, tm1, mu, -1);
      for (int i = 0; i < V; i++) {
        m3_axpy(llderiv[mu][i], naik, m3_mul(M[i], s1b[i]));
        m3_axpy(llderiv[mu][i], naik, sm2[i]);
      }
    }
  }
}
// fat7lDeriv(deriv, gauge, mid, coef, llderiv, llgauge, llmid, naik); ll* may be NULL when naik == 0.
// deriv and llderiv are read, accumulated into and written back.
void orc_fat7l_deriv(void* lat, double* deriv_, const double* gauge_, const double* mid_, const double* coef,
                     double* llderiv_, const double* llgauge_, const double* llmid_, double naik) {
  Lat* l = (Lat*)lat;
  const size_t fs = (size_t)l->vol * 18;
  MF d[4], g[4], m[4], ld[4], lg[4], lm[4];
  for (int k = 0; k < 4; k++) {
    d[k] = mf_load(l, deriv_ + k * fs); g[k] = mf_load(l, gauge_ + k * fs); m[k] = mf_load(l, mid_ + k * fs);
    if (naik != 0.0) { ld[k] = mf_load(l, llderiv_ + k * fs); lg[k] = mf_load(l, llgauge_ + k * fs); lm[k] = mf_load(l, llmid_ + k * fs); }
  }
  fat7l_deriv(l, d, g, m, coef, ld, lg, lm, naik);
  for (int k = 0; k < 4; k++) {
    mf_store(l, deriv_ + k * fs, d[k]);
    if (naik != 0.0) mf_store(l, llderiv_ + k * fs, ld[k]);
  }
}
// HisqCoefs.smearGetForce -> smearedForce(dsdu, dsdsu, dsdsul), hisqsmear.nim:51-84; dsdu accumulates.
void orc_hisq_force(void* lat, const double* u_, const double* c, const double* dsdsu_, const double* dsdsul_, double* dsdu_) {
  Lat* l = (Lat*)lat;
  const size_t fs = (size_t)l->vol * 18, n = fs * 4;
  const int V = l->vol;
  std::vector<double> v(n), w(n);
  orc_makeImpLinks(lat, u_, c, v.data(), nullptr, 0.0, nullptr);
  orc_projectU_field(lat, v.data(), w.data(), 4);
  MF U[4], Vf[4], W[4], su[4], sul[4], f[4], fll[4], t[4], out[4];
  const M3 zero = m3_scale(0.0, m3_load(u_));
  for (int k = 0; k < 4; k++) {
    U[k] = mf_load(l, u_ + k * fs); Vf[k] = mf_load(l, v.data() + k * fs); W[k] = mf_load(l, w.data() + k * fs);
    su[k] = mf_load(l, dsdsu_ + k * fs); sul[k] = mf_load(l, dsdsul_ + k * fs);
    f[k].assign(V, zero); fll[k].assign(V, zero); out[k] = mf_load(l, dsdu_ + k * fs);
  }
  fat7l_deriv(l, f, W, su, c + 5, fll, W, sul, c[10]);                       // asqtadDeriv :16-28
  for (int k = 0; k < 4; k++) {
    t[k].resize(V);
#pragma omp parallel for
    for (int i = 0; i < V; i++) t[k][i] = projectUderiv(W[k][i], Vf[k][i], m3_add(f[k][i], fll[k][i]));   // :77
  }
  fat7l_deriv(l, out, U, t, c, nullptr, nullptr, nullptr, 0.0);              // :78
  for (int k = 0; k < 4; k++) mf_store(l, dsdu_ + k * fs, out[k]);
}

// ---------------------------------------------------------------------------
// Fermion-force outer products (src/physics/stagD.nim:588-664; peqOuter/meqOuter
// src/maths/matrixConcept.nim:717-724: r[i][j] +-= x[i] * conj(y[j])).
// f = nd contiguous link-shaped fields [k][vol][18], hop of field k as in the operator.
// ---------------------------------------------------------------------------
static inline void outer_acc(double* f, const double* a, const double* b, double s) {
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    const double ar = a[2 * i], ai = a[2 * i + 1], br = b[2 * j], bi = b[2 * j + 1];
    f[(i * 3 + j) * 2] += s * (ar * br + ai * bi);
    f[(i * 3 + j) * 2 + 1] += s * (ai * br - ar * bi);
  }
}
// f[k](x) (+)= (x even ? ce : co) * a(x) (x) b(x + h_k)^+ ; accumulate == 0 overwrites
// (the loop of staghmc_spv.nim:833-850 with frc_ind == 0 / > 0)
void orc_outer(void* lat, int nd, double* f, const double* a, const double* b, double ce, double co, int accumulate) {
  Lat* l = (Lat*)lat;
  const size_t fs = (size_t)l->vol * 18;
  for (int k = 0; k < nd; k++) {
    const int mu = nd == 8 ? k / 2 : k, h = (nd == 8 && (k & 1)) ? 1 : 0;
#pragma omp parallel for
    for (int i = 0; i < l->vol; i++) {
      double* fk = f + k * fs + (size_t)i * 18;
      if (!accumulate) for (int e = 0; e < 18; e++) fk[e] = 0.0;
      outer_acc(fk, a + (size_t)i * 6, b + (size_t)l->fwd[mu][h][i] * 6, i < l->half ? ce : co);
    }
  }
}
// stagDeriv, stagD.nim:635-664: even sites +=, odd sites -=, then s.rephase f (nd == 4)
void orc_stagDeriv(void* lat, int nd, double* f, const double* x) {
  orc_outer(lat, nd, f, x, x, 1.0, -1.0, 1);
  if (nd == 4) { orc_setbc(lat, f); orc_stagphase(lat, f); }
}
// stagD2deriv, stagD.nim:588-633: g += c (x) x(+h)^+ ; g -= x (x) c(+h)^+  on all sites
void orc_stagD2deriv(void* lat, int nd, double* g, const double* c, const double* x) {
  orc_outer(lat, nd, g, c, x, 1.0, 1.0, 1);
  orc_outer(lat, nd, g, x, c, -1.0, -1.0, 1);
}

int orc_num_threads() { return omp_get_max_threads(); }
void orc_set_num_threads(int n) { omp_set_num_threads(n); }

}  // extern "C"
