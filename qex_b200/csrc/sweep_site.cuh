// Per-site code of the Dslash sweep -- index arithmetic, link loads (18 / 12 / 14 reals), the 3x3 complex mat-vec, the
// stencil of a site whose hops stay on the rank (stencil_local) or partly leave it (stencil_ghost), the epilogue, and one
// item of the face packing -- together with the argument blocks the kernels take.  Everything here is QB_HD: the kernels
// of dslash.cu are thin wrappers around these functions, and tests/emu/emu_sweep.cpp runs THE SAME functions on the host
// (g++, no GPU) over a lattice split into ranks, against the oracle (tests/test_sweep_emu.py).  Test harness only: the
// library never executes them on the host.
#pragma once
#include "linkops.cuh"   // QB_HD, double2

namespace qb {

#if !defined(__CUDACC__)
inline double rsqrt(double x) { return 1.0 / sqrt(x); }   // host harness only (g++ has no rsqrt)
#endif

constexpr int SWEEP_BLOCK = 128;   // threads per block = sites per tile

// loads with the cache hints the sweep relies on (plain loads on the host)
QB_HD double2 ld_cs(const double2* p) {   // streaming: links are read exactly once per sweep
#if defined(__CUDA_ARCH__)
  return __ldcs(p);
#else
  return *p;
#endif
}
QB_HD double2 ld_nc(const double2* p) {   // read-only path: vectors are re-read by up to 16 neighbours
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}
QB_HD double2 ld_cg(const double2* p) {   // L2 only: ghost data written by a peer GPU must not come from a stale L1 line
#if defined(__CUDA_ARCH__)
  return __ldcg(p);
#else
  return *p;
#endif
}

// A link set as the kernels see it.  Every link class j (0: one-hop, 1: three-hop) is held with
// all 18 reals; a class whose links are s * SU(3) with one |s| for the whole field (naive links
// with staggered phases, Naik links) can ALSO be held as 12 reals: rows 0 and 1, the third row is
// rebuilt in registers as (sigma/|s|) conj(row0 x row1).  sigma = sign of det, one bit per link:
// sgn[par][idx] bit (2 mu + j) for U_{mu,j}(s), bit 8 + (2 mu + j) for U_{mu,j}(s - h_j mu).
struct Links {
  const double2* u;          // [nlinks][2][9][Vh]
  const double2* u12[2];     // per class: [4 mu][2][6][Vh] (12-real) / [4 mu][2][7][Vh] (14-real) or nullptr
  const unsigned short* sgn; // [2][Vh]
  double inv[2];             // 1/|s| per class
};

// Arguments of one parity sweep (kernel parameter block, passed by value).
struct SweepArgs {
  Links L;            // link set
  const double2* x;   // input half field (parity 1-p): x[c*Vh + idx]
  double2* r;         // output half field (parity p)
  const double2* bx;  // half field (parity p) scaled by b, may alias r's parity of the source
  double a, b, sgn, post;   // r = post*(a*r + b*bx + sgn*sum)
  int X[4];
  int Vh;
  int parity;         // output parity p
  int comm_mask;
  // reduction epilogue (DOT): sum over sites of Re<bx, r_new>
  double* partials;
  unsigned int* ticket;
  double* result;
  // != nullptr: the kernel returns at once when *skip != 0 -- the CG's "done" scalar: the host runs a few
  // iterations ahead of the device, and the sweeps it queued past convergence then cost a launch and nothing else
  // on one GPU and on the peer-memory transport (all ranks hold the same value, so every rank skips the same
  // launches).  NOT on the NCCL transport: halo_start has no skip argument, so each of the up to 2 * LAG queued
  // sweeps still runs k_pack and its ncclSend/ncclRecv group before k_sweep / k_exterior return early -- results
  // are unaffected (every rank queues the same exchanges), but an NCCL-transport solve carries those few extra
  // exchanges.
  const double* skip;
};


struct Site {
  int x[4];
  int lex;
};

// idx in [0,Vh) of parity `par` -> local coordinates and lexicographic index
QB_HD Site site_from_idx(int idx, int par, const int* X) {
  Site s;
  int t = idx;
  const int xh = X[0] >> 1;
  int x0h = t % xh; t /= xh;
  s.x[1] = t % X[1]; t /= X[1];
  s.x[2] = t % X[2];
  s.x[3] = t / X[2];
  s.x[0] = 2 * x0h + ((par + s.x[1] + s.x[2] + s.x[3]) & 1);
  s.lex = 2 * idx + (s.x[0] & 1);
  return s;
}

QB_HD int stride_of(int mu, const int* X) {
  return mu == 0 ? 1 : (mu == 1 ? X[0] : (mu == 2 ? X[0] * X[1] : X[0] * X[1] * X[2]));
}

// face index (other three coordinates, lower dimension fastest) >> 1
QB_HD int face_lex(const int* x, int mu, const int* X) {
  int l = 0;
#pragma unroll
  for (int d = 3; d >= 0; d--)
    if (d != mu) l = l * X[d] + x[d];
  return l;
}

// (layer coordinate xmu, face half index, parity) -> full coordinates
QB_HD Site site_from_face(int fidx, int mu, int xmu, int par, const int* X) {
  Site s;
  int d0 = (mu == 0) ? 1 : 0;
  int d1 = (mu <= 1) ? 2 : 1;
  int d2 = (mu <= 2) ? 3 : 2;
  int t = fidx;
  const int eh = X[d0] >> 1;
  int c0h = t % eh; t /= eh;
  int c1 = t % X[d1];
  int c2 = t / X[d1];
  int c0 = 2 * c0h + ((par + xmu + c1 + c2) & 1);
  s.x[mu] = xmu; s.x[d0] = c0; s.x[d1] = c1; s.x[d2] = c2;
  s.lex = ((s.x[3] * X[2] + s.x[2]) * X[1] + s.x[1]) * X[0] + s.x[0];
  return s;
}

// links are read exactly once per sweep: stream them past L1/L2 residency
QB_HD void load_link(double2 (&U)[9], const double2* __restrict__ p, int Vh) {
#pragma unroll
  for (int e = 0; e < 9; e++) U[e] = ld_cs(p + (size_t)e * Vh);
}
// conj(row0 x row1) of the two stored rows
QB_HD void cross_conj(double2 (&c)[3], const double2 (&U)[9]) {
  c[0].x = U[1].x * U[5].x - U[1].y * U[5].y - (U[2].x * U[4].x - U[2].y * U[4].y);
  c[0].y = -(U[1].x * U[5].y + U[1].y * U[5].x - (U[2].x * U[4].y + U[2].y * U[4].x));
  c[1].x = U[2].x * U[3].x - U[2].y * U[3].y - (U[0].x * U[5].x - U[0].y * U[5].y);
  c[1].y = -(U[2].x * U[3].y + U[2].y * U[3].x - (U[0].x * U[5].y + U[0].y * U[5].x));
  c[2].x = U[0].x * U[4].x - U[0].y * U[4].y - (U[1].x * U[3].x - U[1].y * U[3].y);
  c[2].y = -(U[0].x * U[4].y + U[0].y * U[4].x - (U[1].x * U[3].y + U[1].y * U[3].x));
}
// Link U_{mu,j}(par, idx) of a set with NL classes.  RC bit j set: class j is read as 12 reals and
// the third row rebuilt (neg = its sign bit).  RC bit 2+j set: class j is read as 14 reals -- rows 0, 1 and
// one complex factor phi per link, row2 = phi conj(row0 x row1): links that are c * U(3) (projected smeared
// links, Naik products of them), whose determinant phase differs from site to site.
// j is a compile-time constant after unrolling.
template <int NL, int RC>
QB_HD void load_link_k(double2 (&U)[9], const Links& L, int mu, int j, int par,
                                            int idx, int Vh, unsigned int neg) {
  if ((RC >> (2 + j)) & 1) {
    const double2* __restrict__ p = L.u12[j] + ((size_t)(mu * 2 + par) * 7) * Vh + idx;
#pragma unroll
    for (int e = 0; e < 6; e++) U[e] = ld_cs(p + (size_t)e * Vh);
    const double2 phi = ld_cs(p + (size_t)6 * Vh);
    double2 c[3];
    cross_conj(c, U);
#pragma unroll
    for (int e = 0; e < 3; e++) U[6 + e] = make_double2(phi.x * c[e].x - phi.y * c[e].y, phi.x * c[e].y + phi.y * c[e].x);
  } else if ((RC >> j) & 1) {
    const double2* __restrict__ p = L.u12[j] + ((size_t)(mu * 2 + par) * 6) * Vh + idx;
#pragma unroll
    for (int e = 0; e < 6; e++) U[e] = ld_cs(p + (size_t)e * Vh);
    const double f = neg ? -L.inv[j] : L.inv[j];
    // row2 = f * conj(row0 x row1)
    double2 c;
    c.x = U[1].x * U[5].x - U[1].y * U[5].y - (U[2].x * U[4].x - U[2].y * U[4].y);
    c.y = U[1].x * U[5].y + U[1].y * U[5].x - (U[2].x * U[4].y + U[2].y * U[4].x);
    U[6] = make_double2(f * c.x, -f * c.y);
    c.x = U[2].x * U[3].x - U[2].y * U[3].y - (U[0].x * U[5].x - U[0].y * U[5].y);
    c.y = U[2].x * U[3].y + U[2].y * U[3].x - (U[0].x * U[5].y + U[0].y * U[5].x);
    U[7] = make_double2(f * c.x, -f * c.y);
    c.x = U[0].x * U[4].x - U[0].y * U[4].y - (U[1].x * U[3].x - U[1].y * U[3].y);
    c.y = U[0].x * U[4].y + U[0].y * U[4].x - (U[1].x * U[3].y + U[1].y * U[3].x);
    U[8] = make_double2(f * c.x, -f * c.y);
  } else {
    const double2* __restrict__ p = L.u + ((size_t)((mu * NL + j) * 2 + par) * 9) * Vh + idx;
#pragma unroll
    for (int e = 0; e < 9; e++) U[e] = ld_cs(p + (size_t)e * Vh);
  }
}

// vectors are re-read by up to 16 neighbours: keep them cached
QB_HD void load_vec(double2 (&v)[3], const double2* __restrict__ p, int Vh) {
#pragma unroll
  for (int c = 0; c < 3; c++) v[c] = ld_nc(p + (size_t)c * Vh);
}

// acc += U v
QB_HD void mv_add(double2 (&acc)[3], const double2 (&U)[9], const double2 (&v)[3]) {
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int b = 0; b < 3; b++) {
      const double2 u = U[a * 3 + b];
      acc[a].x = fma(u.x, v[b].x, acc[a].x);
      acc[a].x = fma(-u.y, v[b].y, acc[a].x);
      acc[a].y = fma(u.x, v[b].y, acc[a].y);
      acc[a].y = fma(u.y, v[b].x, acc[a].y);
    }
  }
}
// acc -= U^+ v
QB_HD void mv_adj_sub(double2 (&acc)[3], const double2 (&U)[9], const double2 (&v)[3]) {
#pragma unroll
  for (int a = 0; a < 3; a++) {
#pragma unroll
    for (int b = 0; b < 3; b++) {
      const double2 u = U[b * 3 + a];  // conj(u) = (u.x, -u.y)
      acc[a].x = fma(-u.x, v[b].x, acc[a].x);
      acc[a].x = fma(-u.y, v[b].y, acc[a].x);
      acc[a].y = fma(-u.x, v[b].y, acc[a].y);
      acc[a].y = fma(u.y, v[b].x, acc[a].y);
    }
  }
}

QB_HD const double2* link_ptr(const double2* u, int k, int par, int Vh, int idx) {
  return u + ((size_t)(k * 2 + par) * 9) * Vh + idx;
}


// r = post*(a*r + b*bx + sgn*acc) for one site; returns Re<bx, r_new> when DOT
template <bool DOT>
QB_HD double site_epilogue(const SweepArgs& A, int idx, const double2 (&acc)[3]) {
  const int Vh = A.Vh;
  double dot = 0.0;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    double2 o = {A.sgn * acc[c].x, A.sgn * acc[c].y};
    double2 bxv = {0, 0};
    if (A.b != 0.0 || DOT) bxv = A.bx[(size_t)c * Vh + idx];
    if (A.b != 0.0) { o.x = fma(A.b, bxv.x, o.x); o.y = fma(A.b, bxv.y, o.y); }
    if (A.a != 0.0) {
      const double2 ro = A.r[(size_t)c * Vh + idx];
      o.x = fma(A.a, ro.x, o.x); o.y = fma(A.a, ro.y, o.y);
    }
    o.x *= A.post; o.y *= A.post;
    A.r[(size_t)c * Vh + idx] = o;
    if (DOT) dot += bxv.x * o.x + bxv.y * o.y;
  }
  return dot;
}

// sum_k [U_k(s) x(s+h_k) - U_k(s-h_k)^+ x(s-h_k)] over the hops that stay on this rank
template <int NL, int RC>
QB_HD void stencil_local(const SweepArgs& A, const Site& s, int idx, double2 (&acc)[3]) {
  const int Vh = A.Vh, p = A.parity, q = 1 - A.parity;
  const unsigned int sg = (RC & 3) ? A.L.sgn[(size_t)p * Vh + idx] : 0u;
#pragma unroll
  for (int mu = 0; mu < 4; mu++) {
    const int st = stride_of(mu, A.X);
    const int Xm = A.X[mu];
    const bool split = (A.comm_mask >> mu) & 1;
#pragma unroll
    for (int j = 0; j < NL; j++) {
      const int h = (j == 0) ? 1 : 3;
      {  // forward hop: U_k(s) x(s + h mu)
        int y = s.x[mu] + h;
        int ln = s.lex + h * st;
        bool off = y >= Xm;
        if (off) ln -= Xm * st;
        if (!(off && split)) {
          double2 U[9], v[3];
          load_link_k<NL, RC>(U, A.L, mu, j, p, idx, Vh, (sg >> (2 * mu + j)) & 1u);
          load_vec(v, A.x + (ln >> 1), Vh);
          mv_add(acc, U, v);
        }
      }
      {  // backward hop: U_k(s - h mu)^+ x(s - h mu)
        int y = s.x[mu] - h;
        int ln = s.lex - h * st;
        bool off = y < 0;
        if (off) ln += Xm * st;
        if (!(off && split)) {
          double2 U[9], v[3];
          const int ib = ln >> 1;
          load_link_k<NL, RC>(U, A.L, mu, j, q, ib, Vh, (sg >> (8 + 2 * mu + j)) & 1u);
          load_vec(v, A.x + ib, Vh);
          mv_adj_sub(acc, U, v);
        }
      }
    }
  }
}


// ---- 12- / 14-real link storage: what k_compress does for one site (layout.cu) --------------------------------
// Writes rows 0, 1 (and the factor phi of the 14-real form) of every link of site (par, idx) for the classes whose
// mode is 12 or 14, folds the evidence the host decides on into dev / smin / smax (max deviation of the rebuilt third
// row, range of |s|^2) and returns the sign bits of the 12-real form (bit 2 mu + j).
template <int NL>
QB_HD unsigned int compress_site(const double2* __restrict__ u, double2* __restrict__ u12_0, double2* __restrict__ u12_1,
                                 int Vh, int par, int idx, int mode0, int mode1, double (&dev)[2], double (&smin)[2],
                                 double (&smax)[2]) {
  unsigned int bits = 0;
#pragma unroll
  for (int mu = 0; mu < 4; mu++) {
#pragma unroll
    for (int j = 0; j < NL; j++) {
      const int mode = j == 0 ? mode0 : mode1;
      if (mode == 0) continue;
      const int ne = mode == 14 ? 7 : 6;
      const double2* p = u + ((size_t)((mu * NL + j) * 2 + par) * 9) * Vh + idx;
      double2 U[9];
#pragma unroll
      for (int e = 0; e < 9; e++) U[e] = p[(size_t)e * Vh];
      double2* o = (j == 0 ? u12_0 : u12_1) + ((size_t)(mu * 2 + par) * ne) * Vh + idx;
#pragma unroll
      for (int e = 0; e < 6; e++) o[(size_t)e * Vh] = U[e];
      const double s2 = U[0].x * U[0].x + U[0].y * U[0].y + U[1].x * U[1].x + U[1].y * U[1].y +
                        U[2].x * U[2].x + U[2].y * U[2].y;
      const double t2 = U[3].x * U[3].x + U[3].y * U[3].y + U[4].x * U[4].x + U[4].y * U[4].y +
                        U[5].x * U[5].x + U[5].y * U[5].y;
      double2 ov = {0, 0};   // <row0, row1>
#pragma unroll
      for (int e = 0; e < 3; e++) {
        ov.x += U[e].x * U[3 + e].x + U[e].y * U[3 + e].y;
        ov.y += U[e].x * U[3 + e].y - U[e].y * U[3 + e].x;
      }
      double2 c[3];          // conj(row0 x row1)
      c[0].x = U[1].x * U[5].x - U[1].y * U[5].y - (U[2].x * U[4].x - U[2].y * U[4].y);
      c[0].y = -(U[1].x * U[5].y + U[1].y * U[5].x - (U[2].x * U[4].y + U[2].y * U[4].x));
      c[1].x = U[2].x * U[3].x - U[2].y * U[3].y - (U[0].x * U[5].x - U[0].y * U[5].y);
      c[1].y = -(U[2].x * U[3].y + U[2].y * U[3].x - (U[0].x * U[5].y + U[0].y * U[5].x));
      c[2].x = U[0].x * U[4].x - U[0].y * U[4].y - (U[1].x * U[3].x - U[1].y * U[3].y);
      c[2].y = -(U[0].x * U[4].y + U[0].y * U[4].x - (U[1].x * U[3].y + U[1].y * U[3].x));
      const double inv = s2 > 0.0 ? rsqrt(s2) : 0.0;
      double d;
      if (mode == 14) {
        double2 num = {0, 0};   // <c, row2> = sum conj(c_e) row2_e
        double cc = 0.0;
#pragma unroll
        for (int e = 0; e < 3; e++) {
          num.x += c[e].x * U[6 + e].x + c[e].y * U[6 + e].y;
          num.y += c[e].x * U[6 + e].y - c[e].y * U[6 + e].x;
          cc += c[e].x * c[e].x + c[e].y * c[e].y;
        }
        const double icc = cc > 0.0 ? 1.0 / cc : 0.0;
        const double2 phi = {num.x * icc, num.y * icc};
        o[(size_t)6 * Vh] = phi;
        d = 0.0;
#pragma unroll
        for (int e = 0; e < 3; e++) {   // exactly what the sweeps rebuild
          const double rx = phi.x * c[e].x - phi.y * c[e].y, ry = phi.x * c[e].y + phi.y * c[e].x;
          d = fmax(d, fmax(fabs(U[6 + e].x - rx), fabs(U[6 + e].y - ry)) * inv);
        }
        if (!(s2 > 0.0) || !(cc > 0.0)) d = 1.0;
      } else {
        double re = 0.0;       // Re <c, row2> = sigma |c|^2 / |s|
#pragma unroll
        for (int e = 0; e < 3; e++) re += c[e].x * U[6 + e].x + c[e].y * U[6 + e].y;
        const bool neg = re < 0.0;
        if (neg) bits |= 1u << (2 * mu + j);
        const double f = neg ? -inv : inv;
        d = fmax(fabs(t2 - s2), sqrt(ov.x * ov.x + ov.y * ov.y)) * inv * inv;
#pragma unroll
        for (int e = 0; e < 3; e++)
          d = fmax(d, fmax(fabs(U[6 + e].x - f * c[e].x), fabs(U[6 + e].y - f * c[e].y)) * inv);
        if (!(s2 > 0.0)) d = 1.0;
      }
      dev[j] = fmax(dev[j], d);
      smin[j] = fmin(smin[j], s2);
      smax[j] = fmax(smax[j], s2);
    }
  }
  return bits;
}

// high byte of the sign word: the sign bits of the links the BACKWARD hops of site (par, idx) read, U_{mu,j}(s - h_j mu)
template <int NL>
QB_HD unsigned int sign_bwd_site(const unsigned short* sgn, int Vh, int par, int idx, const int* X, unsigned int bits) {
  const Site s = site_from_idx(idx, par, X);   // bits on entry: the site's own low byte
#pragma unroll
  for (int mu = 0; mu < 4; mu++) {
    const int st = stride_of(mu, X);
#pragma unroll
    for (int j = 0; j < NL; j++) {
      const int h = (j == 0) ? 1 : 3;
      int ln = s.lex - h * st;
      if (s.x[mu] - h < 0) ln += X[mu] * st;   // local wrap; off-rank hops never read this bit
      const unsigned int nb = sgn[(size_t)(1 - par) * Vh + (ln >> 1)];
      bits |= ((nb >> (2 * mu + j)) & 1u) << (8 + 2 * mu + j);
    }
  }
  return bits;
}

struct XRank;   // common.cuh

// ---- exterior (NCCL path): one launch for all split directions ------------------------------
struct ExtDir {
  int mu, depth, faceh, nb;
  const double2 *ghost_f, *ghost_b0, *ghost_b1;
  const unsigned int *flag_f, *flag_b;   // p2p: set by the + / - neighbour's pack kernel
};
struct ExtRegion {
  int n[4];      // index counts per direction (n[0] even)
  int type[4];   // 0: whole extent, 1: boundary layers, 2: interior layers
  int end;       // prefix end in threads (sites of one parity)
};
struct ExtArgs {
  Links L;
  double2* r;
  double scale;  // post*sgn
  int X[4], dep[4], nbv[4];
  int Vh, parity, ndir, nreg;
  ExtDir d[4];
  ExtRegion reg[4];
  unsigned int seq;
  int wait;      // p2p: spin on the neighbours' flags before touching the ghosts
  int* err;
  unsigned long long timeout_ns;
  // DOT epilogue: *result = sum over ranks of (*interior + sum over my sites Re<bx, r_new>)
  const double2* bx;
  double* partials;
  unsigned int* ticket;
  double* result;
  const double* interior;
  XRank* xr;
  const double* skip;   // as SweepArgs::skip
};

// thread t of the boundary enumeration -> local coordinates of its site (parity A.parity)
QB_HD void region_site(const ExtArgs& A, int t, int (&x)[4]) {
  int ri = 0, base = 0;
#pragma unroll
  for (int i = 0; i < 3; i++)
    if (i + 1 < A.nreg && t >= A.reg[i].end) { ri = i + 1; base = A.reg[i].end; }
  const ExtRegion& R = A.reg[ri];
  int tt = t - base;
  int id[4];
  const int n0h = R.n[0] >> 1;
  const int i0h = tt % n0h; tt /= n0h;
  id[1] = tt % R.n[1]; tt /= R.n[1];
  id[2] = tt % R.n[2];
  id[3] = tt / R.n[2];
  int par = A.parity;
#pragma unroll
  for (int d = 1; d < 4; d++) par += id[d] + (R.type[d] == 2 ? A.dep[d] : 0);
  par += (R.type[0] == 2 ? A.dep[0] : 0);
  id[0] = 2 * i0h + (par & 1);
#pragma unroll
  for (int d = 0; d < 4; d++) {
    const int i = id[d];
    x[d] = (R.type[d] == 0) ? i : (R.type[d] == 2) ? A.dep[d] + i : (i < A.dep[d] ? i : A.X[d] - A.nbv[d] + i);
  }
}

// One boundary site of the NCCL path's exterior kernel (boundarySB2 of stagD.nim:381-393): thread t of the region
// enumeration adds, for every split direction of its site, the terms whose neighbour lives on another rank -- forward
// hops from the + neighbour's psi layers, backward hops from the products the - neighbour formed.  Returns the site's
// Re<bx, r_new> when DOT.
template <int NL, int RC, bool DOT>
QB_HD double exterior_site(const ExtArgs& A, int t) {
  double dot = 0.0;
  int x[4];
  region_site(A, t, x);
  const int lex = ((x[3] * A.X[2] + x[2]) * A.X[1] + x[1]) * A.X[0] + x[0];
  const int idx = lex >> 1, Vh = A.Vh;
  double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
  const unsigned int sg = (RC & 3) ? A.L.sgn[(size_t)A.parity * Vh + idx] : 0u;
  for (int di = 0; di < A.ndir; di++) {
    const ExtDir& D = A.d[di];
    const int mu = D.mu, Xm = A.X[mu], xmu = x[mu], d = D.depth;
    if (xmu + d < Xm && xmu - d >= 0) continue;   // no off-rank hop in this direction
    const int fidx = face_lex(x, mu, A.X) >> 1;
#pragma unroll
    for (int j = 0; j < NL; j++) {
      const int h = (j == 0) ? 1 : 3;
      if (xmu + h >= Xm) {
        const int l = xmu + h - Xm;
        double2 U[9], v[3];
        load_link_k<NL, RC>(U, A.L, mu, j, A.parity, idx, Vh, (sg >> (2 * mu + j)) & 1u);
#pragma unroll
        for (int c = 0; c < 3; c++) v[c] = ld_cg(D.ghost_f + (size_t)c * d * D.faceh + (size_t)l * D.faceh + fidx);
        mv_add(acc, U, v);
      }
      if (xmu - h < 0) {
        const double2* gb = (j == 0) ? D.ghost_b0 : D.ghost_b1;
#pragma unroll
        for (int c = 0; c < 3; c++) {
          const double2 w = ld_cg(gb + (size_t)c * h * D.faceh + (size_t)xmu * D.faceh + fidx);
          acc[c].x -= w.x; acc[c].y -= w.y;
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; c++) {
    double2 o = A.r[(size_t)c * Vh + idx];
    o.x = fma(A.scale, acc[c].x, o.x); o.y = fma(A.scale, acc[c].y, o.y);
    A.r[(size_t)c * Vh + idx] = o;
    if (DOT) { const double2 bxv = A.bx[(size_t)c * Vh + idx]; dot += bxv.x * o.x + bxv.y * o.y; }
  }
  return dot;
}

// The disjoint boundary regions of fill_ext (dslash.cu): region i = interior in split directions m_1..m_{i-1}, boundary
// layers in m_i, whole extent elsewhere.  Needs E.X, E.ndir, E.d[i].mu and E.nbv filled; returns the number of sites.
inline int plan_regions(ExtArgs& E) {
  int end = 0, nreg = 0;
  for (int i = 0; i < E.ndir; i++) {
    ExtRegion R;
    long long cnt = 1;
    for (int d = 0; d < 4; d++) { R.type[d] = 0; R.n[d] = E.X[d]; }
    for (int j = 0; j < i; j++) { const int m = E.d[j].mu; R.type[m] = 2; R.n[m] = E.X[m] - E.nbv[m]; }
    const int mi = E.d[i].mu;
    R.type[mi] = 1; R.n[mi] = E.nbv[mi];
    for (int d = 0; d < 4; d++) cnt *= R.n[d];
    if (cnt == 0) continue;
    end += (int)(cnt / 2);
    R.end = end;
    E.reg[nreg++] = R;
  }
  E.nreg = nreg;
  return end;
}

// ---- face packing: one work item (see k_pack / k_sweep_mr in dslash.cu) ----------------------
struct PackDir {
  int mu, depth, faceh, end;
  double2 *out_f, *out_b0, *out_b1;
  unsigned int *flag_f, *flag_b;   // flag at the - neighbour (its side 0) / + neighbour (its side 1)
};
struct PackArgs {
  Links L;
  const double2* x;   // half field of parity q (the sweep's input parity)
  int X[4];
  int Vh, q, ndir;
  PackDir d[4];
  unsigned int seq;
  unsigned int* ticket;   // nullptr: no signalling
};

// one work item of the packing (t < A.d[ndir-1].end)
template <int NL, int RC>
QB_HD void pack_item(const PackArgs& A, int t) {
  int di = 0, base = 0;
#pragma unroll
  for (int i = 0; i < 3; i++)
    if (i + 1 < A.ndir && t >= A.d[i].end) { di = i + 1; base = A.d[i].end; }
  const PackDir& P = A.d[di];
  const int tt = t - base;
  const int nf = P.depth * P.faceh;        // forward-psi items
  const int mu = P.mu, Xm = A.X[mu], Vh = A.Vh;
  if (tt < nf) {
    const int l = tt / P.faceh, fidx = tt - l * P.faceh;
    const Site s = site_from_face(fidx, mu, l, A.q, A.X);
    const int idx = s.lex >> 1;
#pragma unroll
    for (int c = 0; c < 3; c++) P.out_f[(size_t)c * nf + tt] = A.x[(size_t)c * Vh + idx];
  } else {
    int tb = tt - nf;
    int j = 0, h = 1;
    if (NL == 2 && tb >= P.faceh) { j = 1; h = 3; tb -= P.faceh; }
    const int jl = tb / P.faceh, fidx = tb - jl * P.faceh;
    const Site s = site_from_face(fidx, mu, Xm - h + jl, A.q, A.X);
    const int idx = s.lex >> 1;
    double2 U[9], v[3], acc[3] = {{0, 0}, {0, 0}, {0, 0}};
    const unsigned int sg = (RC & 3) ? A.L.sgn[(size_t)A.q * Vh + idx] : 0u;
    if (NL == 2 && j == 1) load_link_k<NL, RC>(U, A.L, mu, 1, A.q, idx, Vh, (sg >> (2 * mu + 1)) & 1u);
    else load_link_k<NL, RC>(U, A.L, mu, 0, A.q, idx, Vh, (sg >> (2 * mu)) & 1u);
    load_vec(v, A.x + idx, Vh);
    mv_adj_sub(acc, U, v);
    double2* ob = (j == 0) ? P.out_b0 : P.out_b1;
#pragma unroll
    for (int c = 0; c < 3; c++)
      ob[(size_t)c * h * P.faceh + tb] = make_double2(-acc[c].x, -acc[c].y);
  }
}


struct GhostDir {                        // indexed by mu; entries of unsplit directions stay null
  const double2 *gf, *gb0, *gb1;         // + neighbour's first layers of x ; - neighbour's U_k^+ x per link class
  const unsigned int *flag_f, *flag_b;   // raised by the + / - neighbour's pack blocks
  int faceh;
};
struct MRArgs {
  SweepArgs A;
  PackArgs P;
  GhostDir gd[4];
  const unsigned int* order;   // block -> tile (low 24 bits) | flags it waits for (bit 24 + 2 mu + side)
  int npack, ntiles;
  unsigned int seq;
  int* err;
  unsigned long long timeout_ns;
  XRank* xr;
  unsigned long long* stamps;  // role timestamps of this launch (qexb200_debug_role_times) or nullptr
};
// stamps[0] first block start, [1] last pack block end, [2] last interior tile end, [3] first face tile past its
// wait, [4] last face tile end, [5] longest flag wait, [6] sum of flag waits, [7] face tiles that had to wait
enum { ST_START = 0, ST_PACK_END, ST_INT_END, ST_FACE_START, ST_FACE_END, ST_WAIT_MAX, ST_WAIT_SUM, ST_WAIT_N, ST_N };

template <int NL, int RC>
QB_HD void stencil_ghost(const MRArgs& F, const Site& s, int idx, double2 (&acc)[3]) {
  const SweepArgs& A = F.A;
  const int Vh = A.Vh, p = A.parity, q = 1 - A.parity;
  constexpr int dep = (NL == 2) ? 3 : 1;
  const unsigned int sg = (RC & 3) ? A.L.sgn[(size_t)p * Vh + idx] : 0u;
#pragma unroll
  for (int mu = 0; mu < 4; mu++) {
    const int st = stride_of(mu, A.X);
    const int Xm = A.X[mu], xm = s.x[mu];
    const bool split = (A.comm_mask >> mu) & 1;
    const GhostDir& D = F.gd[mu];
    int fidx = 0;
    if (split && (xm < dep || xm >= Xm - dep)) fidx = face_lex(s.x, mu, A.X) >> 1;
#pragma unroll
    for (int j = 0; j < NL; j++) {
      const int h = (j == 0) ? 1 : 3;
      {  // forward hop: U_k(s) x(s + h mu), x from the + neighbour's first layers when s + h mu left the rank
        const int y = xm + h;
        int ln = s.lex + h * st;
        const bool off = y >= Xm;
        if (off) ln -= Xm * st;
        double2 U[9], v[3];
        load_link_k<NL, RC>(U, A.L, mu, j, p, idx, Vh, (sg >> (2 * mu + j)) & 1u);
        if (off && split) {
#pragma unroll
          for (int c = 0; c < 3; c++)
            v[c] = ld_cg(D.gf + (size_t)c * dep * D.faceh + (size_t)(y - Xm) * D.faceh + fidx);
        } else {
          load_vec(v, A.x + (ln >> 1), Vh);
        }
        mv_add(acc, U, v);
      }
      {  // backward hop: the product U_k^+ x was formed at the source when s - h mu left the rank
        const int y = xm - h;
        int ln = s.lex - h * st;
        const bool off = y < 0;
        if (off) ln += Xm * st;
        if (off && split) {
          const double2* gb = (j == 0) ? D.gb0 : D.gb1;
#pragma unroll
          for (int c = 0; c < 3; c++) {
            const double2 w = ld_cg(gb + (size_t)c * h * D.faceh + (size_t)xm * D.faceh + fidx);
            acc[c].x -= w.x; acc[c].y -= w.y;
          }
        } else {
          double2 U[9], v[3];
          const int ib = ln >> 1;
          load_link_k<NL, RC>(U, A.L, mu, j, q, ib, Vh, (sg >> (8 + 2 * mu + j)) & 1u);
          load_vec(v, A.x + ib, Vh);
          mv_adj_sub(acc, U, v);
        }
      }
    }
  }
}


}  // namespace qb
