// Fermion-force outer products: stagDeriv / stagD2deriv (src/physics/stagD.nim:588-664)
// and the psi (x) psi(x+mu)^+ loop of the HMC force (src/stagg_pv_hmc/staghmc_spv.nim:826-850).
//
// The reference starts nd shifts of the vector, then walks the sites once per direction doing
// peqOuter/meqOuter (src/maths/matrixConcept.nim:717-724) into the link-shaped field.  Here one thread owns
// a site: it loads its own vectors once, gathers the forward neighbours by coordinate arithmetic
// and streams the nd output matrices (read-modify-write only when accumulating).  HBM-bound:
// nd*144 B written (+ nd*144 read when accumulating) against ~(1+nd)*48 B of vector reads per site.
#include "linkops.cuh"
#include "stencil.cuh"

namespace qb {

constexpr int FB = 128;

struct OuterArgs {
  OuterPair p[2];
  double2* f;         // link set [nl*4][2][9][Vh], field k = mu*nl + j, hop 1 (j = 0) or 3 (j = 1)
  int accumulate;
  int X[4];
  int Vh;
  Peers pr;
};
int pool_acquire(double2** out);   // smear.cu: pooled matrix-field temporaries (arena-backed on split lattices)
void pool_release(double2* p);

__device__ __forceinline__ void ld_vec(double2 (&v)[3], const double2* f, const LSite& s, int Vh) {
  const double2* p = f + (size_t)s.par * 3 * Vh + s.idx;
#pragma unroll
  for (int c = 0; c < 3; c++) v[c] = __ldg(p + (size_t)c * Vh);
}

// neighbour vector on a split lattice: from the owner's arena copy over NVLink
__device__ __forceinline__ void ld_vec_peer(double2 (&v)[3], const double2* f, const LSite& s, int Vh, const Peers& pr) {
  const int peer = s.nb == NB_HOME ? pr.me : pr.peer_of[s.nb];
  if (peer == pr.me) { ld_vec(v, f, s, Vh); return; }
  const double2* p = reinterpret_cast<const double2*>(pr.base[peer] + (reinterpret_cast<const char*>(f) - pr.base[pr.me])) +
                     (size_t)s.par * 3 * Vh + s.idx;
#pragma unroll
  for (int c = 0; c < 3; c++) v[c] = ld_peer(p + (size_t)c * Vh);
}

template <int NL, int NP, bool MR>
__global__ void __launch_bounds__(FB) k_outer(const OuterArgs a) {
  const int ds = blockIdx.x * FB + threadIdx.x;
  if (ds >= 2 * a.Vh) return;
  const int par = ds / a.Vh, idx = ds - par * a.Vh;
  const LSite s = lsite(idx, par, a.X);
  double2 va[NP][3];
  double w[NP];
#pragma unroll
  for (int q = 0; q < NP; q++) {
    ld_vec(va[q], a.p[q].a, s, a.Vh);
    w[q] = par ? a.p[q].co : a.p[q].ce;
  }
#pragma unroll
  for (int mu = 0; mu < 4; mu++) {
#pragma unroll
    for (int j = 0; j < NL; j++) {
      const LSite t = lhop(s, mu, j ? 3 : 1, a.X);
      double2* fp = a.f + ((size_t)((mu * NL + j) * 2 + par) * 9) * a.Vh + idx;
      M3 r;
      if (a.accumulate) {
#pragma unroll
        for (int e = 0; e < 9; e++) r.e[e] = __ldcs(fp + (size_t)e * a.Vh);
      } else {
        r = m3_zero();
      }
#pragma unroll
      for (int q = 0; q < NP; q++) {
        double2 vb[3];
        if (MR) ld_vec_peer(vb, a.p[q].b, t, a.Vh, a.pr);
        else ld_vec(vb, a.p[q].b, t, a.Vh);
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
          for (int k = 0; k < 3; k++) {   // x[i] * conj(y[k])
            r.e[i * 3 + k].x += w[q] * (va[q][i].x * vb[k].x + va[q][i].y * vb[k].y);
            r.e[i * 3 + k].y += w[q] * (va[q][i].y * vb[k].x - va[q][i].x * vb[k].y);
          }
      }
#pragma unroll
      for (int e = 0; e < 9; e++) __stcs(fp + (size_t)e * a.Vh, r.e[e]);
    }
  }
}

// f[k](x) (+)= sum_q w_q(x) a_q(x) (x) b_q(x + h_k)^+, np = 1 or 2 pairs.  On a split lattice the shifted
// operands are first copied into arena temporaries (vectors live in ordinary allocations), a cross-rank
// barrier follows, and the kernel reads the off-rank neighbours from the owners' copies.
int outer_products(Gauge& f, const OuterPair* pairs, int np, bool accumulate) {
  Context& c = ctx();
  if (np < 1 || np > 2) return fail("outer_products: 1 or 2 pairs");
  const int nl = f.nlinks / 4;
  for (int d = 0; d < 4; d++)
    if (c.g.X[d] < (nl == 2 ? 4 : 2)) return fail("outer_products: local extent %d too small", c.g.X[d]);
  OuterArgs a = {};
  for (int q = 0; q < np; q++) a.p[q] = pairs[q];
  a.f = f.d; a.accumulate = accumulate ? 1 : 0; a.Vh = c.g.Vh;
  for (int d = 0; d < 4; d++) a.X[d] = c.g.X[d];
  const int grid = (c.g.V + FB - 1) / FB;
  const bool mr = c.g.nranks > 1;
  double2* tmp[2] = {nullptr, nullptr};
  if (mr) {
    QB_CHECK(peers_fill(&a.pr));
    for (int q = 0; q < np; q++) {
      if (q == 1 && pairs[1].b == pairs[0].b) { a.p[1].b = a.p[0].b; continue; }
      QB_CHECK(pool_acquire(&tmp[q]));
      if (!in_arena(tmp[q])) return fail("force outer products on a split lattice need the peer arena (QEXB200_ARENA_FIELDS > 0)");
      QB_CUDA(cudaMemcpyAsync(tmp[q], pairs[q].b, sizeof(double2) * 6 * (size_t)c.g.Vh, cudaMemcpyDeviceToDevice, c.stream));
      a.p[q].b = tmp[q];
    }
    QB_CHECK(xbarrier());
  }
#define QB_OUTER(MRV)                                                              \
  do {                                                                             \
    if (nl == 1 && np == 1) k_outer<1, 1, MRV><<<grid, FB, 0, c.stream>>>(a);      \
    else if (nl == 1) k_outer<1, 2, MRV><<<grid, FB, 0, c.stream>>>(a);            \
    else if (np == 1) k_outer<2, 1, MRV><<<grid, FB, 0, c.stream>>>(a);            \
    else k_outer<2, 2, MRV><<<grid, FB, 0, c.stream>>>(a);                         \
  } while (0)
  if (mr) QB_OUTER(true); else QB_OUTER(false);
#undef QB_OUTER
  count_launch();
  QB_CUDA(cudaGetLastError());
  if (mr) {
    QB_CHECK(xbarrier());   // the peers have finished reading the temporaries
    for (int q = 0; q < 2; q++) pool_release(tmp[q]);
  }
  return 0;
}

}  // namespace qb
