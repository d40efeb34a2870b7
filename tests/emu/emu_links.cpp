// TEST HARNESS ONLY: runs the per-site functions of qex_b200/csrc/linkops.cuh on the host (they are
// __host__ __device__) over a whole field, so the coordinate arithmetic and the 3x3 algebra of the
// device link pipeline can be checked against the oracle on a machine without a GPU.  Compiled by
// tests/test_links_emu.py with g++; never part of libqex_b200.so.
#include "../../qex_b200/csrc/linkops.cuh"

using namespace qb;

#include <vector>
extern "C" {
// fields in the DEVICE layout double2[2][9][Vh]
void emu_staple_field(const double* G, const double* L, int dG, int dL, const int* X, double* out) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      const LSite s = lsite(idx, par, X);
      const M3 st = staple_pair((const double2*)G, (const double2*)L, dG, dL, s, X, Vh);
      m3_st((double2*)out, s, Vh, st);
    }
}
void emu_project_field(const double* in, const int* X, double* out) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      const LSite s = lsite(idx, par, X);
      m3_st((double2*)out, s, Vh, project_u(m3_ld((const double2*)in, s, Vh)));
    }
}
// One rank of a split lattice: arenas = the nranks per-rank buffers (same layout on every rank), fields given as
// byte offsets into the arena; a site's off-rank neighbours are read from the owner's arena (m3_ldx<true>).
void emu_staple_field_mr(char* const* arenas, int nranks, int me, const unsigned char* peer_of, long long offG,
                         long long offL, int dG, int dL, const int* X, double* out) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  Peers pr;
  for (int r = 0; r < nranks; r++) pr.base[r] = arenas[r];
  for (int k = 0; k < 81; k++) pr.peer_of[k] = peer_of[k];
  pr.me = me;
  const double2* G = (const double2*)(arenas[me] + offG);
  const double2* L = (const double2*)(arenas[me] + offL);
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      const LSite s = lsite(idx, par, X);
      m3_st((double2*)out, s, Vh, staple_pair_x<true>(G, L, dG, dL, s, X, Vh, pr));
    }
}
// out = projectUderiv(x, chain) site by site
void emu_project_u_deriv_field(const double* x, const double* chain, const int* X, double* out) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      const LSite s = lsite(idx, par, X);
      m3_st((double2*)out, s, Vh, project_u_deriv(m3_ld((const double2*)x, s, Vh), m3_ld((const double2*)chain, s, Vh)));
    }
}
// f1, f2 += symStapleDeriv(g1 (d1), g2 (d2), c)
void emu_staple_deriv_field(double* f1, double* f2, const double* g1, int d1, const double* g2, int d2, const double* c,
                            const int* X) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  Peers none; none.me = 0;
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      const LSite s = lsite(idx, par, X);
      M3 a = m3_ld((const double2*)f1, s, Vh), b = m3_ld((const double2*)f2, s, Vh);
      staple_deriv_x<false>(a, b, (const double2*)g1, d1, (const double2*)g2, d2, (const double2*)c, s, X, Vh, none);
      m3_st((double2*)f1, s, Vh, a);
      m3_st((double2*)f2, s, Vh, b);
    }
}
// Host ring for staple_deriv_pipe_thread.  late = 0: a request is delivered when it is made (the earliest a
// cp.async can land: a slot refilled before its last read would show).  late = 1: requests are queued per
// commit group and delivered only when a wait demands it (the latest they may land: a read ahead of its wait,
// or a wrong wait count, would return stale data).
struct HostRing {
  struct Req { double2* dst; const double2* src; int stride; };
  double2 buf[2 * 27];
  std::vector<std::vector<Req>> groups;
  std::vector<Req> open;
  int late;
  explicit HostRing(int l) : late(l) { for (auto& b : buf) b = make_double2(1e300, -1e300); }
  static void deliver(const Req& r) { for (int e = 0; e < 9; e++) r.dst[e] = r.src[(size_t)e * r.stride]; }
  void put(int slot, int m, const double2* f, const LSite& s, int Vh) {
    Req r = {buf + (slot * 3 + m) * 9, f + (size_t)s.par * 9 * Vh + s.idx, Vh};
    if (late) open.push_back(r); else deliver(r);
  }
  M3 get(int slot, int m) const {
    M3 r;
    for (int e = 0; e < 9; e++) r.e[e] = buf[(slot * 3 + m) * 9 + e];
    return r;
  }
  void commit() { groups.push_back(open); open.clear(); }
  void drain(size_t keep) {
    while (groups.size() > keep) { for (const Req& r : groups.front()) deliver(r); groups.erase(groups.begin()); }
  }
  void wait1() { drain(1); }
  void wait0() { drain(0); }
};
// f1, f2 += symStapleDeriv through the ring pipeline: nblocks persistent blocks of pb threads, run one thread after another
void emu_staple_deriv_pipe_field(double* f1, double* f2, const double* g1, int d1, const double* g2, int d2, const double* c,
                                 const int* X, int nblocks, int pb, int late) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  const int ntiles = 2 * ((Vh + pb - 1) / pb);
  for (int b = 0; b < nblocks; b++)
    for (int t = 0; t < pb; t++) {
      HostRing ring(late);
      staple_deriv_pipe_thread(ring, (double2*)f1, (double2*)f2, (const double2*)g1, d1, (const double2*)g2, d2,
                               (const double2*)c, X, Vh, ntiles, b, nblocks, t, pb);
    }
}
// out += naik * naik-derivative(U, M) in direction mu
void emu_naik_deriv_field(double* out, const double* U, const double* M, int mu, double naik, const int* X) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  Peers none; none.me = 0;
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      const LSite s = lsite(idx, par, X);
      M3 o = m3_ld((const double2*)out, s, Vh);
      m3_axpy(o, naik, naik_deriv_x<false>((const double2*)U, (const double2*)M, mu, s, X, Vh, none));
      m3_st((double2*)out, s, Vh, o);
    }
}
void emu_project_tah_field(const double* in, const int* X, double* out) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      const LSite s = lsite(idx, par, X);
      m3_st((double2*)out, s, Vh, project_tah(m3_ld((const double2*)in, s, Vh)));
    }
}
// coordinates of (par, idx) and the index of its neighbour: out[6] = x0..x3, par', idx'
void emu_hop(int idx, int par, int mu, int d, const int* X, int* out) {
  const LSite s = lsite(idx, par, X);
  const LSite t = lhop(s, mu, d, X);
  for (int k = 0; k < 4; k++) out[k] = s.x[k];
  out[4] = t.par; out[5] = t.idx;
}
}
