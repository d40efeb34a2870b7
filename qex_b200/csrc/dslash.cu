// Staggered Dslash parity sweep for sm_100a (HBM-bound fp64 stencil, no tensor cores).
//
// Restates WHAT src/physics/stagD.nim:348-395 (stagD2) computes,
//   r[s] = a r[s] + b x[s] + sum_k [ U_k(s) x(s+h_k) - U_k(s-h_k)^+ x(s-h_k) ],  s of parity p,
// with the 3-phase shape of the reference (start shifts / local / boundary,
// stagD.nim:355-393) mapped to: face-pack kernels + NCCL exchange on a second
// stream, an interior kernel that skips off-rank neighbours, and per-direction
// exterior kernels that add the halo terms.
//
// One thread per output site; all loads are 128-bit (one complex double) and
// coalesced along the site index.  Links are streamed with ld.global.cs, vectors
// go through the read-only path so the 8/16 neighbour re-reads hit L1/L2.
#include "stencil.cuh"

namespace qb {

constexpr int SWEEP_BLOCK = 128;

template <int NL, bool DOT>
__global__ void __launch_bounds__(SWEEP_BLOCK) k_sweep(const SweepArgs A) {
  const int idx = blockIdx.x * SWEEP_BLOCK + threadIdx.x;
  double dot = 0.0;
  if (idx < A.Vh) {
    const int Vh = A.Vh, p = A.parity, q = 1 - A.parity;
    const Site s = site_from_idx(idx, p, A.X);
    double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
#pragma unroll
    for (int mu = 0; mu < 4; mu++) {
      const int st = stride_of(mu, A.X);
      const int Xm = A.X[mu];
      const bool split = (A.comm_mask >> mu) & 1;
#pragma unroll
      for (int j = 0; j < NL; j++) {
        const int h = (j == 0) ? 1 : 3;
        const int k = mu * NL + j;
        {  // forward hop: U_k(s) x(s + h mu)
          int y = s.x[mu] + h;
          int ln = s.lex + h * st;
          bool off = y >= Xm;
          if (off) ln -= Xm * st;
          if (!(off && split)) {
            double2 U[9], v[3];
            load_link(U, link_ptr(A.u, k, p, Vh, idx), Vh);
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
            load_link(U, link_ptr(A.u, k, q, Vh, ib), Vh);
            load_vec(v, A.x + ib, Vh);
            mv_adj_sub(acc, U, v);
          }
        }
      }
    }
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
  }
  if (DOT) grid_reduce<SWEEP_BLOCK>(dot, A.partials, A.ticket, A.result);
}

// Exterior kernel of split direction mu: adds the terms whose neighbour lives on
// another rank (boundarySB2 of stagD.nim:381-393).  ghost_f holds the + neighbour's
// first `depth` layers of x; ghost_b[j] holds U_k^+ x of the - neighbour's last h_j layers.
struct ExtArgs {
  const double2* u;
  double2* r;
  const double2* ghost_f;
  const double2* ghost_b[2];
  double scale;  // post*sgn
  int X[4];
  int Vh, parity, mu, nb, depth, faceh;
};

template <int NL>
__global__ void __launch_bounds__(SWEEP_BLOCK) k_exterior(const ExtArgs A) {
  const int t = blockIdx.x * SWEEP_BLOCK + threadIdx.x;
  if (t >= A.nb * A.faceh) return;
  const int b = t / A.faceh, fidx = t - b * A.faceh;
  const int mu = A.mu, Xm = A.X[mu], d = A.depth;
  const int xmu = (b < d) ? b : Xm - A.nb + b;
  const Site s = site_from_face(fidx, mu, xmu, A.parity, A.X);
  const int idx = s.lex >> 1, Vh = A.Vh;
  double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
#pragma unroll
  for (int j = 0; j < NL; j++) {
    const int h = (j == 0) ? 1 : 3;
    const int k = mu * NL + j;
    if (xmu + h >= Xm) {
      const int l = xmu + h - Xm;
      double2 U[9], v[3];
      load_link(U, link_ptr(A.u, k, A.parity, Vh, idx), Vh);
#pragma unroll
      for (int c = 0; c < 3; c++) v[c] = A.ghost_f[(size_t)c * d * A.faceh + (size_t)l * A.faceh + fidx];
      mv_add(acc, U, v);
    }
    if (xmu - h < 0) {
#pragma unroll
      for (int c = 0; c < 3; c++) {
        const double2 w = A.ghost_b[j][(size_t)c * h * A.faceh + (size_t)xmu * A.faceh + fidx];
        acc[c].x -= w.x; acc[c].y -= w.y;
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; c++) {
    double2 o = A.r[(size_t)c * Vh + idx];
    o.x = fma(A.scale, acc[c].x, o.x); o.y = fma(A.scale, acc[c].y, o.y);
    A.r[(size_t)c * Vh + idx] = o;
  }
}

// Face packing (startSB of stagD.nim:355-361): x of my first `depth` layers for the
// - neighbour, and U_k^+ x of my last h_j layers for the + neighbour.
struct PackArgs {
  const double2* u;
  const double2* x;   // half field of parity q (the sweep's input parity)
  double2* out_f;
  double2* out_b[2];
  int X[4];
  int Vh, q, mu, depth, faceh;
};

template <int NL>
__global__ void __launch_bounds__(SWEEP_BLOCK) k_pack(const PackArgs A) {
  const int t = blockIdx.x * SWEEP_BLOCK + threadIdx.x;
  const int nf = A.depth * A.faceh;        // forward-psi items
  const int ntot = nf + ((NL == 2) ? 4 : 1) * A.faceh;
  if (t >= ntot) return;
  const int mu = A.mu, Xm = A.X[mu], Vh = A.Vh;
  if (t < nf) {
    const int l = t / A.faceh, fidx = t - l * A.faceh;
    const Site s = site_from_face(fidx, mu, l, A.q, A.X);
    const int idx = s.lex >> 1;
#pragma unroll
    for (int c = 0; c < 3; c++) A.out_f[(size_t)c * nf + t] = A.x[(size_t)c * Vh + idx];
  } else {
    int tb = t - nf;
    int j = 0, h = 1;
    if (NL == 2 && tb >= A.faceh) { j = 1; h = 3; tb -= A.faceh; }
    const int jl = tb / A.faceh, fidx = tb - jl * A.faceh;
    const Site s = site_from_face(fidx, mu, Xm - h + jl, A.q, A.X);
    const int idx = s.lex >> 1;
    double2 U[9], v[3], acc[3] = {{0, 0}, {0, 0}, {0, 0}};
    load_link(U, link_ptr(A.u, mu * NL + j, A.q, Vh, idx), Vh);
    load_vec(v, A.x + idx, Vh);
    mv_adj_sub(acc, U, v);
#pragma unroll
    for (int c = 0; c < 3; c++)
      A.out_b[j][(size_t)c * h * A.faceh + tb] = make_double2(-acc[c].x, -acc[c].y);
  }
}

int pack_faces(const Gauge& g, int q, const double2* x) {
  Context& c = ctx();
  for (const HaloDir& hd : c.halo) {
    PackArgs P;
    P.u = g.d; P.x = x;
    P.out_f = c.sendbuf + hd.off_f;
    P.out_b[0] = c.sendbuf + hd.off_b;
    P.out_b[1] = c.sendbuf + hd.off_b + (size_t)3 * hd.faceh;
    for (int i = 0; i < 4; i++) P.X[i] = c.g.X[i];
    P.Vh = c.g.Vh; P.q = q; P.mu = hd.mu; P.depth = hd.depth; P.faceh = hd.faceh;
    const int nl = g.nlinks / 4;
    const int ntot = (hd.depth + (nl == 2 ? 4 : 1)) * hd.faceh;
    const int grid = (ntot + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
    if (nl == 2) k_pack<2><<<grid, SWEEP_BLOCK, 0, c.stream>>>(P);
    else k_pack<1><<<grid, SWEEP_BLOCK, 0, c.stream>>>(P);
    count_launch();
  }
  QB_CUDA(cudaGetLastError());
  return 0;
}

static int exterior(const Gauge& g, int parity, double2* r, double scale) {
  Context& c = ctx();
  for (const HaloDir& hd : c.halo) {
    ExtArgs E;
    E.u = g.d; E.r = r;
    E.ghost_f = c.recvbuf + hd.off_f;
    E.ghost_b[0] = c.recvbuf + hd.off_b;
    E.ghost_b[1] = c.recvbuf + hd.off_b + (size_t)3 * hd.faceh;
    E.scale = scale;
    for (int i = 0; i < 4; i++) E.X[i] = c.g.X[i];
    E.Vh = c.g.Vh; E.parity = parity; E.mu = hd.mu; E.depth = hd.depth; E.faceh = hd.faceh;
    const int Xm = c.g.X[hd.mu];
    E.nb = (2 * hd.depth < Xm) ? 2 * hd.depth : Xm;
    const int n = E.nb * hd.faceh;
    const int grid = (n + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
    if (g.nlinks == 8) k_exterior<2><<<grid, SWEEP_BLOCK, 0, c.stream>>>(E);
    else k_exterior<1><<<grid, SWEEP_BLOCK, 0, c.stream>>>(E);
    count_launch();
  }
  QB_CUDA(cudaGetLastError());
  return 0;
}

int sweep(const Gauge& g, int parity, double2* r, const double2* x, const double2* bx, double a,
          double b, double sgn, double post, double* dot_out) {
  Context& c = ctx();
  if (!g.loaded) return fail("sweep: gauge not loaded");
  const bool multi = c.g.comm_mask != 0;
  if (multi) QB_CHECK(halo_start(g, parity, x));

  SweepArgs A;
  A.u = g.d; A.x = x; A.r = r; A.bx = bx ? bx : r;
  A.a = a; A.b = bx ? b : 0.0; A.sgn = sgn; A.post = post;
  for (int i = 0; i < 4; i++) A.X[i] = c.g.X[i];
  A.Vh = c.g.Vh; A.parity = parity; A.comm_mask = c.g.comm_mask;
  A.partials = c.d_partials; A.ticket = c.d_ticket; A.result = dot_out;
  const int grid = (c.g.Vh + SWEEP_BLOCK - 1) / SWEEP_BLOCK;
  const bool fused_dot = dot_out && !multi;
  if (g.nlinks == 8) {
    if (fused_dot) k_sweep<2, true><<<grid, SWEEP_BLOCK, 0, c.stream>>>(A);
    else k_sweep<2, false><<<grid, SWEEP_BLOCK, 0, c.stream>>>(A);
  } else {
    if (fused_dot) k_sweep<1, true><<<grid, SWEEP_BLOCK, 0, c.stream>>>(A);
    else k_sweep<1, false><<<grid, SWEEP_BLOCK, 0, c.stream>>>(A);
  }
  count_launch();
  QB_CUDA(cudaGetLastError());
  if (multi) {
    QB_CHECK(halo_wait());
    QB_CHECK(exterior(g, parity, r, post * sgn));
    if (dot_out) QB_CHECK(blas_redot(bx, r, (size_t)3 * c.g.Vh, dot_out));
  }
  return 0;
}

}  // namespace qb
