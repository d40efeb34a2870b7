// Per-site arithmetic of the device link pipeline: 3x3 complex products, the staple
// of one (side link, middle link) pair gathered straight from the checkerboarded fields, projectU.
// Everything here is QB_HD so that tests/emu_links.cpp can run the very same site functions on the
// host against the oracle before GPU time is spent (test harness only; the library never does).
//
// A "matrix field" is double2 f[2 parity][9 (row,col)][Vh]; a link set stores nlinks of them.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#if defined(__CUDACC__)
#define QB_HD __host__ __device__ __forceinline__
#else
#define QB_HD inline
#endif

namespace qb {

struct M3 {
  double2 e[9];
};

struct LSite {   // local coordinates + parity + checkerboard index
  int x[4];
  int par, idx;
  int nb;        // rank displacement of the site: sum_d (disp_d + 1) 3^d, NB_HOME = this rank
};
constexpr int NB_HOME = 40;

// Peer view of the symmetric arena (multi-rank link pipeline): every rank allocates its link-shaped fields
// at the same offsets of an IPC-mapped arena, so the neighbour's copy of field f is at
// base[peer] + (f - base[me]).  peer_of[nb] = rank owning displacement nb (periodic in the rank grid).
struct Peers {
  char* base[16];
  unsigned char peer_of[81];
  int me;
};

QB_HD LSite lsite(int idx, int par, const int* X) {
  LSite s;
  int t = idx;
  const int xh = X[0] >> 1;
  const int x0h = t % xh; t /= xh;
  s.x[1] = t % X[1]; t /= X[1];
  s.x[2] = t % X[2];
  s.x[3] = t / X[2];
  s.x[0] = 2 * x0h + ((par + s.x[1] + s.x[2] + s.x[3]) & 1);
  s.par = par; s.idx = idx; s.nb = NB_HOME;
  return s;
}
// one hop of d = +-1 (or +-k) in direction mu with local wrap-around (single rank: periodic)
// (selects instead of s.x[mu]: a run-time index would put the coordinates in local memory)
QB_HD LSite lhop(LSite s, int mu, int d, const int* X) {
  const int xm = mu == 0 ? s.x[0] : (mu == 1 ? s.x[1] : (mu == 2 ? s.x[2] : s.x[3]));
  const int Xm = mu == 0 ? X[0] : (mu == 1 ? X[1] : (mu == 2 ? X[2] : X[3]));
  int c = xm + d;
  const int p3 = mu == 0 ? 1 : (mu == 1 ? 3 : (mu == 2 ? 9 : 27));
  if (c >= Xm) { c -= Xm; s.nb += p3; }
  if (c < 0) { c += Xm; s.nb -= p3; }
  s.x[0] = mu == 0 ? c : s.x[0];
  s.x[1] = mu == 1 ? c : s.x[1];
  s.x[2] = mu == 2 ? c : s.x[2];
  s.x[3] = mu == 3 ? c : s.x[3];
  if (d & 1) s.par ^= 1;
  s.idx = (((s.x[3] * X[2] + s.x[2]) * X[1] + s.x[1]) * X[0] + s.x[0]) >> 1;
  return s;
}

QB_HD M3 m3_ld(const double2* f, const LSite& s, int Vh) {
  M3 m;
  const double2* p = f + (size_t)s.par * 9 * Vh + s.idx;
#pragma unroll
  for (int e = 0; e < 9; e++) m.e[e] = p[(size_t)e * Vh];
  return m;
}
// 16-byte load of peer memory: system scope, never served from a stale L1 line
QB_HD double2 ld_peer(const double2* p) {
#if defined(__CUDA_ARCH__)
  double2 v;
  asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
#else
  return *p;
#endif
}
// MR: a site that left the rank (s.nb != NB_HOME) is read from the owner's arena over NVLink
template <bool MR>
QB_HD M3 m3_ldx(const double2* f, const LSite& s, int Vh, const Peers& pr) {
  if (!MR) return m3_ld(f, s, Vh);
  const int peer = s.nb == NB_HOME ? pr.me : pr.peer_of[s.nb];
  if (peer == pr.me) return m3_ld(f, s, Vh);
  const double2* p = reinterpret_cast<const double2*>(pr.base[peer] + (reinterpret_cast<const char*>(f) - pr.base[pr.me])) +
                     (size_t)s.par * 9 * Vh + s.idx;
  M3 m;
#pragma unroll
  for (int e = 0; e < 9; e++) m.e[e] = ld_peer(p + (size_t)e * Vh);
  return m;
}
QB_HD void m3_st(double2* f, const LSite& s, int Vh, const M3& m) {
  double2* p = f + (size_t)s.par * 9 * Vh + s.idx;
#pragma unroll
  for (int e = 0; e < 9; e++) p[(size_t)e * Vh] = m.e[e];
}

QB_HD double2 c_mul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// a * b
QB_HD M3 m3_mul(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      double re = 0.0, im = 0.0;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const double2 p = a.e[i * 3 + k], q = b.e[k * 3 + j];   // 4 FMAs per complex multiply-add
        re = fma(p.x, q.x, re); re = fma(-p.y, q.y, re);
        im = fma(p.x, q.y, im); im = fma(p.y, q.x, im);
      }
      r.e[i * 3 + j] = make_double2(re, im);
    }
  return r;
}
// a * b^+
QB_HD M3 m3_mul_na(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      double re = 0.0, im = 0.0;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const double2 p = a.e[i * 3 + k], q = b.e[j * 3 + k];   // conj(q)
        re = fma(p.x, q.x, re); re = fma(p.y, q.y, re);
        im = fma(p.y, q.x, im); im = fma(-p.x, q.y, im);
      }
      r.e[i * 3 + j] = make_double2(re, im);
    }
  return r;
}
// a^+ * b
QB_HD M3 m3_mul_an(const M3& a, const M3& b) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      double re = 0.0, im = 0.0;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const double2 p = a.e[k * 3 + i], q = b.e[k * 3 + j];   // conj(p)
        re = fma(p.x, q.x, re); re = fma(p.y, q.y, re);
        im = fma(p.x, q.y, im); im = fma(-p.y, q.x, im);
      }
      r.e[i * 3 + j] = make_double2(re, im);
    }
  return r;
}
QB_HD void m3_axpy(M3& y, double a, const M3& x) {
#pragma unroll
  for (int e = 0; e < 9; e++) { y.e[e].x += a * x.e[e].x; y.e[e].y += a * x.e[e].y; }
}
QB_HD M3 m3_zero() {
  M3 r;
#pragma unroll
  for (int e = 0; e < 9; e++) r.e[e] = make_double2(0.0, 0.0);
  return r;
}

// The two staples of middle link L (direction dL) with side link G (direction dG) at site s:
//   G(x) L(x+dG) G(x+dL)^+  +  G(x-dG)^+ L(x-dG) G(x-dG+dL)
// = computeGenStaple's staple (src/gauge/fat7l.nim:24-76, gauge[nu] = G, link = L, mu = dL, nu = dG)
// = symStaple's bracket (src/gauge/smearutil.nim:3-17, g1 = G, g2 = L).
// The reference forms the lower staple at x-dG and shifts it; here it is gathered in place.
template <bool MR>
QB_HD M3 staple_pair_x(const double2* G, const double2* L, int dG, int dL, const LSite& s, const int* X, int Vh,
                       const Peers& pr) {
  const LSite sG = lhop(s, dG, +1, X);
  const LSite sL = lhop(s, dL, +1, X);
  M3 r = m3_mul_na(m3_mul(m3_ld(G, s, Vh), m3_ldx<MR>(L, sG, Vh, pr)), m3_ldx<MR>(G, sL, Vh, pr));
  const LSite sb = lhop(s, dG, -1, X);
  const LSite sbL = lhop(sb, dL, +1, X);
  const M3 lo = m3_mul(m3_mul_an(m3_ldx<MR>(G, sb, Vh, pr), m3_ldx<MR>(L, sb, Vh, pr)), m3_ldx<MR>(G, sbL, Vh, pr));
  m3_axpy(r, 1.0, lo);
  return r;
}
QB_HD M3 staple_pair(const double2* G, const double2* L, int dG, int dL, const LSite& s, const int* X, int Vh) {
  Peers none;
  none.me = 0;
  return staple_pair_x<false>(G, L, dG, dL, s, X, Vh, none);
}

// ---- projectU (src/maths/matrixFunctions.nim:79-181,301-312) --------------------------------
// eigenvalues of a Hermitian 3x3 from tr, tr(x^2), det: eigs3, :79-111
QB_HD void eigs3(double& e0, double& e1, double& e2, double tr, double p2, double det) {
  const double tr3 = (1.0 / 3.0) * tr;
  const double p23 = (1.0 / 3.0) * p2;
  const double tr32 = tr3 * tr3;
  const double q = fabs(0.5 * (p23 - tr32));
  const double r = 0.25 * tr3 * (5 * tr32 - p2) - 0.5 * det;
  const double sq = sqrt(q);
  const double sq3 = q * sq;
  const double isq3 = 1.0 / sq3;
  const double isq3c = fmin(3e38, fmax(-3e38, isq3));
  const double rsq3c = r * isq3c;
  const double rsq3 = fmin(1.0, fmax(-1.0, rsq3c));
  const double t = (1.0 / 3.0) * acos(rsq3);
  const double st = sin(t), ct = cos(t);
  const double sqc = sq * ct;
  const double sqs = 1.73205080756887729352 * sq * st;
  const double ll = tr3 + sqc;
  e0 = tr3 - 2 * sqc; e1 = ll + sqs; e2 = ll - sqs;
}
QB_HD double m3_det_re(const M3& x) {
  const double2 a = c_mul(x.e[0], x.e[4]), b = c_mul(x.e[1], x.e[3]);
  const double2 c = c_mul(x.e[2], x.e[3]), d = c_mul(x.e[0], x.e[5]);
  const double2 f = c_mul(x.e[1], x.e[5]), g = c_mul(x.e[2], x.e[4]);
  const double2 t0 = c_mul(make_double2(a.x - b.x, a.y - b.y), x.e[8]);
  const double2 t1 = c_mul(make_double2(c.x - d.x, c.y - d.y), x.e[7]);
  const double2 t2 = c_mul(make_double2(f.x - g.x, f.y - g.y), x.e[6]);
  return t0.x + t1.x + t2.x;
}
// (x^+ x + 1e-20)^(-1/2): projectUrsqrt :301-306 through rsqrtPHM3 :126-181
QB_HD M3 project_u_rsqrt(const M3& x) {
  M3 h = m3_mul_an(x, x);
  h.e[0].x += 1e-20; h.e[4].x += 1e-20; h.e[8].x += 1e-20;
  const double tr = h.e[0].x + h.e[4].x + h.e[8].x;
  const M3 h2 = m3_mul(h, h);
  const double p2 = h2.e[0].x + h2.e[4].x + h2.e[8].x;
  const double det = m3_det_re(h);
  double l0, l1, l2;
  eigs3(l0, l1, l2, tr, p2, det);
  const double sl0 = sqrt(fabs(l0)), sl1 = sqrt(fabs(l1)), sl2 = sqrt(fabs(l2));
  const double u = sl0 + sl1 + sl2;
  const double w = sl0 * sl1 * sl2;
  const double d = w * (sl0 + sl1) * (sl0 + sl2) * (sl1 + sl2);
  const double di = 1.0 / d;
  const double c0 = (w * u * u + l0 * sl0 * (l1 + l2) + l1 * sl1 * (l0 + l2) + l2 * sl2 * (l0 + l1)) * di;
  const double c1 = -(tr * u + w) * di;
  const double c2 = u * di;
  M3 rs;
#pragma unroll
  for (int e = 0; e < 9; e++) rs.e[e] = make_double2(c1 * h.e[e].x + c2 * h2.e[e].x, c1 * h.e[e].y + c2 * h2.e[e].y);
  rs.e[0].x += c0; rs.e[4].x += c0; rs.e[8].x += c0;
  return rs;
}
// x (x^+ x + 1e-20)^(-1/2): projectU :308-312
QB_HD M3 project_u(const M3& x) { return m3_mul(x, project_u_rsqrt(x)); }

// ---- projectUderiv (src/maths/matrixFunctions.nim:321-357) ----------------------------------------
QB_HD double2 c_sub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }
QB_HD double2 c_add(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
QB_HD double2 c_inv(double2 a) { const double n = 1.0 / (a.x * a.x + a.y * a.y); return make_double2(a.x * n, -a.y * n); }
// adjugate, src/maths/projUderiv.nim:17-35
QB_HD M3 m3_adjugate(const M3& x) {
  M3 r;
  r.e[0] = c_sub(c_mul(x.e[4], x.e[8]), c_mul(x.e[5], x.e[7]));
  r.e[1] = c_sub(c_mul(x.e[7], x.e[2]), c_mul(x.e[8], x.e[1]));
  r.e[2] = c_sub(c_mul(x.e[1], x.e[5]), c_mul(x.e[2], x.e[4]));
  r.e[3] = c_sub(c_mul(x.e[5], x.e[6]), c_mul(x.e[3], x.e[8]));
  r.e[4] = c_sub(c_mul(x.e[8], x.e[0]), c_mul(x.e[6], x.e[2]));
  r.e[5] = c_sub(c_mul(x.e[2], x.e[3]), c_mul(x.e[0], x.e[5]));
  r.e[6] = c_sub(c_mul(x.e[3], x.e[7]), c_mul(x.e[4], x.e[6]));
  r.e[7] = c_sub(c_mul(x.e[6], x.e[1]), c_mul(x.e[7], x.e[0]));
  r.e[8] = c_sub(c_mul(x.e[0], x.e[4]), c_mul(x.e[1], x.e[3]));
  return r;
}
// inverse, src/maths/matinv.nim:89-113 (adjugate / determinant)
QB_HD M3 m3_inverse(const M3& x) {
  const M3 ad = m3_adjugate(x);
  const double2 det = c_add(c_add(c_mul(ad.e[8], x.e[8]), c_mul(ad.e[5], x.e[7])), c_mul(ad.e[2], x.e[6]));
  const double2 idet = c_inv(det);
  M3 r;
#pragma unroll
  for (int e = 0; e < 9; e++) r.e[e] = c_mul(idet, ad.e[e]);
  return r;
}
// A X + X A = C for X, src/maths/projUderiv.nim:122-146
QB_HD M3 m3_sylsolve(const M3& a, const M3& c) {
  const M3 ad = m3_adjugate(a);
  const double2 t = c_add(c_add(a.e[0], a.e[4]), a.e[8]);
  const double2 s = c_add(c_add(ad.e[0], ad.e[4]), ad.e[8]);
  const double2 r = c_add(c_add(c_mul(a.e[0], ad.e[0]), c_mul(a.e[1], ad.e[3])), c_mul(a.e[2], ad.e[6]));
  const M3 ac = m3_mul(a, c), ca = m3_mul(c, a), aca = m3_mul(ac, a), adc = m3_mul(ad, c), cad = m3_mul(c, ad),
           adcad = m3_mul(adc, ad);
  const double2 st = c_sub(c_mul(s, t), r);
  const double2 c2 = c_inv(make_double2(2.0 * st.x, 2.0 * st.y));
  const double2 c0 = c_mul(c2, c_add(s, c_mul(t, t)));
  const double2 c1 = c_mul(c2, c_mul(t, c_inv(r)));
  const double2 c4 = c_mul(c2, t);
  M3 x;
#pragma unroll
  for (int e = 0; e < 9; e++) {
    const double2 v = c_add(c_add(c_mul(c0, c.e[e]), c_mul(c1, adcad.e[e])),
                            c_mul(c2, c_sub(c_sub(aca.e[e], adc.e[e]), cad.e[e])));
    x.e[e] = c_sub(v, c_mul(c4, c_add(ac.e[e], ca.e[e])));
  }
  return x;
}
// projectUderiv(r, u, x, chain) :329-351 with u = x z formed from the same z (the 3-argument form :353-357)
QB_HD M3 project_u_deriv(const M3& x, const M3& chain) {
  const M3 z = project_u_rsqrt(x);
  const M3 u = m3_mul(x, z);
  const M3 y = m3_inverse(z);
  M3 r = m3_mul(chain, z);
  const M3 t1 = m3_mul_an(u, r);
  const M3 t2 = m3_sylsolve(y, t1);
  M3 t3;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++)
      t3.e[i * 3 + j] = make_double2(t2.e[i * 3 + j].x + t2.e[j * 3 + i].x, t2.e[i * 3 + j].y - t2.e[j * 3 + i].y);
  m3_axpy(r, -1.0, m3_mul(x, t3));
  return r;
}

// ---- symStapleDeriv (src/gauge/smearutil.nim:24-54), gathered in place --------------------------------
// g1: side link (direction d1), g2: middle link (d2), c: chain of the smeared link (direction d2).
//   f1(x) += g2(x) g1(x+d2) c(x+d1)^+ + c(x) g1(x+d2) g2(x+d1)^+ + [g2^+ g1 c(+d1) + c^+ g1 g2(+d1)](x-d2)
//   f2(x) += g1(x) c(x+d1) g1(x+d2)^+ + [g1^+ c g1(+d2)](x-d1)
template <bool MR>
QB_HD void staple_deriv_x(M3& f1, M3& f2, const double2* g1, int d1, const double2* g2, int d2, const double2* c,
                          const LSite& s, const int* X, int Vh, const Peers& pr) {
  const LSite s1 = lhop(s, d1, +1, X), s2 = lhop(s, d2, +1, X);
  const M3 G1x2 = m3_ldx<MR>(g1, s2, Vh, pr), Cx1 = m3_ldx<MR>(c, s1, Vh, pr);
  {
    const M3 t = m3_mul_na(m3_mul(m3_ld(g2, s, Vh), G1x2), Cx1);
    m3_axpy(f1, 1.0, t);
    const M3 v = m3_mul_na(m3_mul(m3_ld(c, s, Vh), G1x2), m3_ldx<MR>(g2, s1, Vh, pr));
    m3_axpy(f1, 1.0, v);
    const M3 w = m3_mul_na(m3_mul(m3_ld(g1, s, Vh), Cx1), G1x2);
    m3_axpy(f2, 1.0, w);
  }
  {
    const LSite y = lhop(s, d2, -1, X), y1 = lhop(y, d1, +1, X);
    const M3 G1y = m3_ldx<MR>(g1, y, Vh, pr);
    const M3 a = m3_mul(m3_mul_an(m3_ldx<MR>(g2, y, Vh, pr), G1y), m3_ldx<MR>(c, y1, Vh, pr));
    m3_axpy(f1, 1.0, a);
    const M3 b = m3_mul(m3_mul_an(m3_ldx<MR>(c, y, Vh, pr), G1y), m3_ldx<MR>(g2, y1, Vh, pr));
    m3_axpy(f1, 1.0, b);
  }
  {
    const LSite y = lhop(s, d1, -1, X), y2 = lhop(y, d2, +1, X);
    const M3 a = m3_mul(m3_mul_an(m3_ldx<MR>(g1, y, Vh, pr), m3_ldx<MR>(c, y, Vh, pr)), m3_ldx<MR>(g1, y2, Vh, pr));
    m3_axpy(f2, 1.0, a);
  }
}
// ---- the same derivative with the 14 gathers streamed through a two-slot ring ---------------------------
// (smear.cu k_staple_deriv_pipe; experimental, QEXB200_DERIV_PIPE=1).  The 14 matrices of a site are consumed in
// five half-steps of <= 3 matrices; half-step n is read from slot n & 1, which is refilled at once with the
// matrices of half-step n + 2 (of this site or of the thread's next site), so every request has two half-steps
// of products between issue and use.  G1(x+d2), C(x+d1) and G1(x-d2) stay in registers across half-steps.
// R is the ring policy: put(slot, m, field, site, Vh) requests a matrix into position m of a slot, get(slot, m)
// reads it, commit() closes a group of requests, wait1() returns once all groups but the newest have landed,
// wait0() once all have.  Device: cp.async into shared memory.  tests/emu: a host ring that delivers either at
// once or as late as the waits allow, which checks the slot reuse order and the wait counts without a GPU.
// Arithmetic and summation order are those of staple_deriv_x.
template <class R>
QB_HD void deriv_ring_fill(R& ring, int slot, int k, const double2* g1, int d1, const double2* g2, int d2,
                           const double2* c, const LSite& s, const int* X, int Vh) {
  if (k == 0) {
    ring.put(slot, 0, g2, s, Vh); ring.put(slot, 1, g1, lhop(s, d2, +1, X), Vh); ring.put(slot, 2, c, lhop(s, d1, +1, X), Vh);
  } else if (k == 1) {
    ring.put(slot, 0, c, s, Vh); ring.put(slot, 1, g2, lhop(s, d1, +1, X), Vh); ring.put(slot, 2, g1, s, Vh);
  } else if (k == 2) {
    const LSite y = lhop(s, d2, -1, X);
    ring.put(slot, 0, g2, y, Vh); ring.put(slot, 1, g1, y, Vh); ring.put(slot, 2, c, lhop(y, d1, +1, X), Vh);
  } else if (k == 3) {
    const LSite y = lhop(s, d2, -1, X);
    ring.put(slot, 0, c, y, Vh); ring.put(slot, 1, g2, lhop(y, d1, +1, X), Vh);
  } else {
    const LSite z = lhop(s, d1, -1, X);
    ring.put(slot, 0, g1, z, Vh); ring.put(slot, 1, c, z, Vh); ring.put(slot, 2, g1, lhop(z, d2, +1, X), Vh);
  }
}
// One thread of a persistent block: tiles block, block + nblocks, ... (nblocks even: the parity never changes),
// site (tile >> 1) * pb + tid of each.  F1, F2 accumulate (read-modify-write of the thread's own site only).
template <class R>
QB_HD void staple_deriv_pipe_thread(R& ring, double2* F1, double2* F2, const double2* g1, int d1, const double2* g2,
                                    int d2, const double2* c, const int* X, int Vh, int ntiles, int block,
                                    int nblocks, int tid, int pb) {
  const int par = block & 1;
  int tile = block;
  int idx = (tile >> 1) * pb + tid;
  bool live = tile < ntiles && idx < Vh;
  LSite s = lsite(live ? idx : 0, par, X);
  if (live) deriv_ring_fill(ring, 0, 0, g1, d1, g2, d2, c, s, X, Vh);
  ring.commit();
  if (live) deriv_ring_fill(ring, 1, 1, g1, d1, g2, d2, c, s, X, Vh);
  ring.commit();
  int p = 0;   // slot of this site's half-step 0 (five half-steps per site: it alternates)
  while (tile < ntiles) {
    const int ntile = tile + nblocks;
    const int nidx = (ntile >> 1) * pb + tid;
    const bool nlive = ntile < ntiles && nidx < Vh;
    const LSite ns = lsite(nlive ? nidx : 0, par, X);
    const int q = p ^ 1;
    M3 f1 = m3_zero(), f2 = m3_zero();
    if (live) { f1 = m3_ld(F1, s, Vh); f2 = m3_ld(F2, s, Vh); }
    M3 G1x2, Cx1, G1y;
    ring.wait1();   // half-step 0: g2(x), g1(x+d2), c(x+d1)
    {
      const M3 A = ring.get(p, 0);
      G1x2 = ring.get(p, 1); Cx1 = ring.get(p, 2);
      if (live) deriv_ring_fill(ring, p, 2, g1, d1, g2, d2, c, s, X, Vh);
      ring.commit();
      m3_axpy(f1, 1.0, m3_mul_na(m3_mul(A, G1x2), Cx1));
    }
    ring.wait1();   // half-step 1: c(x), g2(x+d1), g1(x)
    {
      const M3 A = ring.get(q, 0), B = ring.get(q, 1), G1s = ring.get(q, 2);
      if (live) deriv_ring_fill(ring, q, 3, g1, d1, g2, d2, c, s, X, Vh);
      ring.commit();
      m3_axpy(f1, 1.0, m3_mul_na(m3_mul(A, G1x2), B));
      m3_axpy(f2, 1.0, m3_mul_na(m3_mul(G1s, Cx1), G1x2));
    }
    ring.wait1();   // half-step 2: g2(y), g1(y), c(y+d1), y = x-d2
    {
      const M3 A = ring.get(p, 0);
      G1y = ring.get(p, 1);
      const M3 B = ring.get(p, 2);
      if (live) deriv_ring_fill(ring, p, 4, g1, d1, g2, d2, c, s, X, Vh);
      ring.commit();
      m3_axpy(f1, 1.0, m3_mul(m3_mul_an(A, G1y), B));
    }
    ring.wait1();   // half-step 3: c(y), g2(y+d1)
    {
      const M3 A = ring.get(q, 0), B = ring.get(q, 1);
      if (nlive) deriv_ring_fill(ring, q, 0, g1, d1, g2, d2, c, ns, X, Vh);
      ring.commit();
      m3_axpy(f1, 1.0, m3_mul(m3_mul_an(A, G1y), B));
    }
    ring.wait1();   // half-step 4: g1(z), c(z), g1(z+d2), z = x-d1
    {
      const M3 A = ring.get(p, 0), B = ring.get(p, 1), C = ring.get(p, 2);
      if (nlive) deriv_ring_fill(ring, p, 1, g1, d1, g2, d2, c, ns, X, Vh);
      ring.commit();
      m3_axpy(f2, 1.0, m3_mul(m3_mul_an(A, B), C));
    }
    if (live) { m3_st(F1, s, Vh, f1); m3_st(F2, s, Vh, f2); }
    tile = ntile; idx = nidx; live = nlive; s = ns; p = q;
  }
  ring.wait0();
}

// Naik term of fat7lDeriv (src/gauge/fat7lderiv.nim:300-326), gathered in place; U = llgauge[mu], M = llmid[mu]:
//   (U^+ M)(x-mu) U(x+mu)^+  +  M(x) (U(x+mu) U(x+2mu))^+  +  U(x-mu)^+ U(x-2mu)^+ M(x-2mu)
template <bool MR>
QB_HD M3 naik_deriv_x(const double2* U, const double2* M, int mu, const LSite& s, const int* X, int Vh, const Peers& pr) {
  const LSite sp = lhop(s, mu, +1, X), spp = lhop(sp, mu, +1, X), sm = lhop(s, mu, -1, X), smm = lhop(sm, mu, -1, X);
  const M3 Um = m3_ldx<MR>(U, sm, Vh, pr), Up = m3_ldx<MR>(U, sp, Vh, pr);
  M3 r = m3_mul_na(m3_mul_an(Um, m3_ldx<MR>(M, sm, Vh, pr)), Up);
  m3_axpy(r, 1.0, m3_mul_na(m3_ld(M, s, Vh), m3_mul(Up, m3_ldx<MR>(U, spp, Vh, pr))));
  m3_axpy(r, 1.0, m3_mul_an(Um, m3_mul_an(m3_ldx<MR>(U, smm, Vh, pr), m3_ldx<MR>(M, smm, Vh, pr))));
  return r;
}
// projectTAH, src/maths/matrixFunctions.nim:375-380: (x - x^+)/2 minus its trace / 3
QB_HD M3 project_tah(const M3& x) {
  M3 r;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++)
      r.e[i * 3 + j] = make_double2(0.5 * (x.e[i * 3 + j].x - x.e[j * 3 + i].x), 0.5 * (x.e[i * 3 + j].y + x.e[j * 3 + i].y));
  const double d = (r.e[0].y + r.e[4].y + r.e[8].y) * (1.0 / 3.0);   // the trace is purely imaginary
  r.e[0].y -= d; r.e[4].y -= d; r.e[8].y -= d;
  r.e[0].x = 0.0; r.e[4].x = 0.0; r.e[8].x = 0.0;
  return r;
}

}  // namespace qb
