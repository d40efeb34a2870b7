// TEST HARNESS ONLY: one parity sweep of the staggered Dslash on a lattice split over several ranks, with every rank
// emulated in this process and the per-site functions of qex_b200/csrc/sweep_site.cuh -- the very code the kernels
// k_pack / k_sweep_mr of dslash.cu wrap -- executed on the host: pack_item writes the psi layers and the U^+ psi
// products into the neighbours' receive buffers, then every tile of the block -> tile table runs stencil_local or
// stencil_ghost and the epilogue.  What the GPU adds on top (flags, fences, the grid) is ordering only.  Compiled by
// tests/test_sweep_emu.py with g++; never part of libqex_b200.so.
#include "../../qex_b200/csrc/sweep_site.cuh"

using namespace qb;

struct EmuHalo {   // one split direction of one rank: the fields of qexb200_halo_dir the sweep uses
  int mu, depth, faceh, rank_plus, rank_minus;
  long long off_f, off_b;
};

struct EmuRecon {   // per rank: the compressed copies k_compress / k_sign_bwd leave (layout.cu compress_links)
  const double* u12[2];
  const unsigned short* sgn;
  double inv[2];
};

template <int NL, int RC>
static void sweep_all(int nranks, const int* X, int comm_mask, int parity, const double* const* links, const EmuRecon* rec,
                      const double* const* x, double* const* r, double a, double b, double sgn, double post,
                      const EmuHalo* halo, int ndir, double* const* recv, const unsigned int* const* tiles, int ntiles) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  const int q = 1 - parity;
  // phase 1 -- what the pack blocks do (fill_pack of dslash.cu, peer-memory destinations)
  for (int me = 0; me < nranks; me++) {
    PackArgs P;
    P.L.u = (const double2*)links[me];
    for (int j = 0; j < 2; j++) { P.L.u12[j] = rec ? (const double2*)rec[me].u12[j] : nullptr; P.L.inv[j] = rec ? rec[me].inv[j] : 1.0; }
    P.L.sgn = rec ? rec[me].sgn : nullptr;
    P.x = (const double2*)x[me] + (size_t)q * 3 * Vh;
    for (int d = 0; d < 4; d++) P.X[d] = X[d];
    P.Vh = Vh; P.q = q; P.ndir = ndir; P.seq = 1; P.ticket = nullptr;
    int end = 0;
    for (int i = 0; i < ndir; i++) {
      const EmuHalo& h = halo[me * 4 + i];
      PackDir& D = P.d[i];
      D.mu = h.mu; D.depth = h.depth; D.faceh = h.faceh;
      end += (h.depth + (NL == 2 ? 4 : 1)) * h.faceh;
      D.end = end;
      D.out_f = (double2*)recv[h.rank_minus] + h.off_f;
      D.out_b0 = (double2*)recv[h.rank_plus] + h.off_b;
      D.out_b1 = D.out_b0 + (size_t)3 * h.faceh;
      D.flag_f = D.flag_b = nullptr;
    }
    for (int t = 0; t < end; t++) pack_item<NL, RC>(P, t);
  }
  // phase 2 -- what the tile blocks do, in table order (interior tiles first, face tiles last)
  for (int me = 0; me < nranks; me++) {
    MRArgs F;
    SweepArgs& A = F.A;
    A.L = Links();
    A.L.u = (const double2*)links[me];
    for (int j = 0; j < 2; j++) { A.L.u12[j] = rec ? (const double2*)rec[me].u12[j] : nullptr; A.L.inv[j] = rec ? rec[me].inv[j] : 1.0; }
    A.L.sgn = rec ? rec[me].sgn : nullptr;
    A.x = (const double2*)x[me] + (size_t)q * 3 * Vh;
    A.r = (double2*)r[me] + (size_t)parity * 3 * Vh;
    A.bx = (const double2*)x[me] + (size_t)parity * 3 * Vh;
    A.a = a; A.b = b; A.sgn = sgn; A.post = post;
    for (int d = 0; d < 4; d++) A.X[d] = X[d];
    A.Vh = Vh; A.parity = parity; A.comm_mask = comm_mask;
    A.partials = nullptr; A.ticket = nullptr; A.result = nullptr; A.skip = nullptr;
    for (int mu = 0; mu < 4; mu++) F.gd[mu] = GhostDir{nullptr, nullptr, nullptr, nullptr, nullptr, 0};
    for (int i = 0; i < ndir; i++) {
      const EmuHalo& h = halo[me * 4 + i];
      GhostDir& D = F.gd[h.mu];
      D.gf = (const double2*)recv[me] + h.off_f;
      D.gb0 = (const double2*)recv[me] + h.off_b;
      D.gb1 = D.gb0 + (size_t)3 * h.faceh;
      D.faceh = h.faceh;
    }
    for (int bi = 0; bi < ntiles; bi++) {
      const unsigned int e = tiles[me][bi];
      const int tile = (int)(e & 0xffffffu);
      const unsigned int need = e >> 24;
      for (int tid = 0; tid < SWEEP_BLOCK; tid++) {
        const int idx = tile * SWEEP_BLOCK + tid;
        if (idx >= Vh) continue;
        const Site s = site_from_idx(idx, parity, A.X);
        double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
        if (need) stencil_ghost<NL, RC>(F, s, idx, acc);
        else stencil_local<NL, RC>(A, s, idx, acc);
        site_epilogue<false>(A, idx, acc);
      }
    }
  }
}

// The NCCL transport's three phases (pack -> send/recv -> interior k_sweep -> k_exterior) with the same site functions:
// pack_item into each rank's own send buffer, the exchange of comm.cu halo_start as plain copies (psi faces come from
// the + neighbour, product faces from the - neighbour), stencil_local on every site (off-rank hops skipped), then
// exterior_site over the disjoint boundary regions of plan_regions.
template <int NL, int RC>
static void sweep_nccl(int nranks, const int* X, int comm_mask, int parity, const double* const* links, const EmuRecon* rec,
                       const double* const* x, double* const* r, double a, double b, double sgn, double post,
                       const EmuHalo* halo, int ndir, double* const* send, double* const* recv) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  const int q = 1 - parity;
  auto links_of = [&](int me) {
    Links L = Links();
    L.u = (const double2*)links[me];
    for (int j = 0; j < 2; j++) { L.u12[j] = rec ? (const double2*)rec[me].u12[j] : nullptr; L.inv[j] = rec ? rec[me].inv[j] : 1.0; }
    L.sgn = rec ? rec[me].sgn : nullptr;
    return L;
  };
  for (int me = 0; me < nranks; me++) {
    PackArgs P;
    P.L = links_of(me);
    P.x = (const double2*)x[me] + (size_t)q * 3 * Vh;
    for (int d = 0; d < 4; d++) P.X[d] = X[d];
    P.Vh = Vh; P.q = q; P.ndir = ndir; P.seq = 1; P.ticket = nullptr;
    int end = 0;
    for (int i = 0; i < ndir; i++) {
      const EmuHalo& h = halo[me * 4 + i];
      PackDir& D = P.d[i];
      D.mu = h.mu; D.depth = h.depth; D.faceh = h.faceh;
      end += (h.depth + (NL == 2 ? 4 : 1)) * h.faceh;
      D.end = end;
      D.out_f = (double2*)send[me] + h.off_f;
      D.out_b0 = (double2*)send[me] + h.off_b;
      D.out_b1 = D.out_b0 + (size_t)3 * h.faceh;
      D.flag_f = D.flag_b = nullptr;
    }
    for (int t = 0; t < end; t++) pack_item<NL, RC>(P, t);
  }
  for (int me = 0; me < nranks; me++)       // ncclSend / ncclRecv pairs of halo_start
    for (int i = 0; i < ndir; i++) {
      const EmuHalo& h = halo[me * 4 + i];
      const size_t nf = (size_t)3 * h.depth * h.faceh, nb = (size_t)3 * (NL == 2 ? 4 : 1) * h.faceh;
      for (size_t k = 0; k < 2 * nf; k++) recv[me][2 * h.off_f + k] = send[h.rank_plus][2 * h.off_f + k];
      for (size_t k = 0; k < 2 * nb; k++) recv[me][2 * h.off_b + k] = send[h.rank_minus][2 * h.off_b + k];
    }
  for (int me = 0; me < nranks; me++) {
    SweepArgs A;
    A.L = links_of(me);
    A.x = (const double2*)x[me] + (size_t)q * 3 * Vh;
    A.r = (double2*)r[me] + (size_t)parity * 3 * Vh;
    A.bx = (const double2*)x[me] + (size_t)parity * 3 * Vh;
    A.a = a; A.b = b; A.sgn = sgn; A.post = post;
    for (int d = 0; d < 4; d++) A.X[d] = X[d];
    A.Vh = Vh; A.parity = parity; A.comm_mask = comm_mask;
    A.partials = nullptr; A.ticket = nullptr; A.result = nullptr; A.skip = nullptr;
    for (int idx = 0; idx < Vh; idx++) {    // interior kernel: every site, hops that stay on the rank
      const Site s = site_from_idx(idx, parity, A.X);
      double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
      stencil_local<NL, RC>(A, s, idx, acc);
      site_epilogue<false>(A, idx, acc);
    }
    ExtArgs E;                              // fill_ext of dslash.cu
    E.L = A.L; E.r = A.r; E.scale = post * sgn; E.bx = A.bx;
    E.partials = nullptr; E.ticket = nullptr; E.result = nullptr; E.interior = nullptr; E.xr = nullptr; E.skip = nullptr;
    for (int d = 0; d < 4; d++) { E.X[d] = X[d]; E.dep[d] = 0; E.nbv[d] = 0; }
    E.Vh = Vh; E.parity = parity; E.ndir = ndir; E.seq = 1; E.wait = 0; E.err = nullptr; E.timeout_ns = 0;
    for (int i = 0; i < ndir; i++) {
      const EmuHalo& h = halo[me * 4 + i];
      ExtDir& D = E.d[i];
      D.mu = h.mu; D.depth = h.depth; D.faceh = h.faceh;
      D.nb = (2 * h.depth < X[h.mu]) ? 2 * h.depth : X[h.mu];
      D.ghost_f = (const double2*)recv[me] + h.off_f;
      D.ghost_b0 = (const double2*)recv[me] + h.off_b;
      D.ghost_b1 = D.ghost_b0 + (size_t)3 * h.faceh;
      D.flag_f = D.flag_b = nullptr;
      E.dep[h.mu] = h.depth; E.nbv[h.mu] = D.nb;
    }
    const int end = plan_regions(E);
    for (int t = 0; t < end; t++) exterior_site<NL, RC, false>(E, t);
  }
}

extern "C" {
// fields in the DEVICE layout: links[r] = double2[nlinks][2][9][Vh], x[r] / r[r] = double2[2][3][Vh];
// recv[r] = the receive buffer of rank r; halo[r * 4 + i], i < ndir; tiles[r] = qexb200_plan_tiles of rank r;
// rc = the storage bits of Gauge::rc() (bit j: class j as 12 reals, bit 2 + j: as 14 reals), rec = the compressed copies
int emu_sweep_mr(int nranks, int nl, int rc, const int* X, int comm_mask, int parity, const double* const* links,
                 const EmuRecon* rec, const double* const* x, double* const* r, double a, double b, double sgn, double post,
                 const EmuHalo* halo, int ndir, double* const* recv, const unsigned int* const* tiles, int ntiles) {
#define EMU_CASE(NLV, RCV)                                                                                              \
  if (nl == NLV && rc == RCV) {                                                                                         \
    sweep_all<NLV, RCV>(nranks, X, comm_mask, parity, links, RCV ? rec : nullptr, x, r, a, b, sgn, post, halo, ndir, recv, \
                        tiles, ntiles);                                                                                 \
    return 0;                                                                                                           \
  }
  EMU_CASE(1, 0) EMU_CASE(1, 1) EMU_CASE(1, 4)
  EMU_CASE(2, 0) EMU_CASE(2, 2) EMU_CASE(2, 3) EMU_CASE(2, 8) EMU_CASE(2, 12) EMU_CASE(2, 9) EMU_CASE(2, 6)
#undef EMU_CASE
  return 1;
}
int emu_sweep_nccl(int nranks, int nl, int rc, const int* X, int comm_mask, int parity, const double* const* links,
                   const EmuRecon* rec, const double* const* x, double* const* r, double a, double b, double sgn, double post,
                   const EmuHalo* halo, int ndir, double* const* send, double* const* recv) {
#define EMU_CASE(NLV, RCV)                                                                                                     \
  if (nl == NLV && rc == RCV) {                                                                                                \
    sweep_nccl<NLV, RCV>(nranks, X, comm_mask, parity, links, RCV ? rec : nullptr, x, r, a, b, sgn, post, halo, ndir, send, recv); \
    return 0;                                                                                                                  \
  }
  EMU_CASE(1, 0) EMU_CASE(1, 1) EMU_CASE(1, 4)
  EMU_CASE(2, 0) EMU_CASE(2, 2) EMU_CASE(2, 3) EMU_CASE(2, 8) EMU_CASE(2, 12) EMU_CASE(2, 9) EMU_CASE(2, 6)
#undef EMU_CASE
  return 1;
}
// ---- one rank at a time: for the world_size-2/4 gloo runs of tests/multi_worker.py --cpu-sweep, where every process
// is one rank and the send/recv buffers travel through torch.distributed (18-real links) --------------------------------
static Links plain_links(const double* links) {
  Links L = Links();
  L.u = (const double2*)links; L.u12[0] = L.u12[1] = nullptr; L.sgn = nullptr; L.inv[0] = L.inv[1] = 1.0;
  return L;
}
// phase 1 of the NCCL transport on this rank: pack the faces of x (parity 1 - parity) into `send`
int emu_rank_pack(int nl, const int* X, int parity, const double* links, const double* x, const EmuHalo* halo, int ndir,
                  double* send) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2, q = 1 - parity;
  PackArgs P;
  P.L = plain_links(links);
  P.x = (const double2*)x + (size_t)q * 3 * Vh;
  for (int d = 0; d < 4; d++) P.X[d] = X[d];
  P.Vh = Vh; P.q = q; P.ndir = ndir; P.seq = 1; P.ticket = nullptr;
  int end = 0;
  for (int i = 0; i < ndir; i++) {
    PackDir& D = P.d[i];
    D.mu = halo[i].mu; D.depth = halo[i].depth; D.faceh = halo[i].faceh;
    end += (halo[i].depth + (nl == 2 ? 4 : 1)) * halo[i].faceh;
    D.end = end;
    D.out_f = (double2*)send + halo[i].off_f;
    D.out_b0 = (double2*)send + halo[i].off_b;
    D.out_b1 = D.out_b0 + (size_t)3 * halo[i].faceh;
    D.flag_f = D.flag_b = nullptr;
  }
  for (int t = 0; t < end; t++) { if (nl == 2) pack_item<2, 0>(P, t); else pack_item<1, 0>(P, t); }
  return 0;
}
// phases 2 + 3 on this rank: interior sweep, then the boundary terms from the received buffer `recv`
int emu_rank_sweep(int nl, const int* X, int comm_mask, int parity, const double* links, const double* x, double* r,
                   double a, double b, double sgn, double post, const EmuHalo* halo, int ndir, const double* recv) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2, q = 1 - parity;
  SweepArgs A;
  A.L = plain_links(links);
  A.x = (const double2*)x + (size_t)q * 3 * Vh;
  A.r = (double2*)r + (size_t)parity * 3 * Vh;
  A.bx = (const double2*)x + (size_t)parity * 3 * Vh;
  A.a = a; A.b = b; A.sgn = sgn; A.post = post;
  for (int d = 0; d < 4; d++) A.X[d] = X[d];
  A.Vh = Vh; A.parity = parity; A.comm_mask = comm_mask;
  A.partials = nullptr; A.ticket = nullptr; A.result = nullptr; A.skip = nullptr;
  for (int idx = 0; idx < Vh; idx++) {
    const Site s = site_from_idx(idx, parity, A.X);
    double2 acc[3] = {{0, 0}, {0, 0}, {0, 0}};
    if (nl == 2) stencil_local<2, 0>(A, s, idx, acc); else stencil_local<1, 0>(A, s, idx, acc);
    site_epilogue<false>(A, idx, acc);
  }
  ExtArgs E;
  E.L = A.L; E.r = A.r; E.scale = post * sgn; E.bx = A.bx;
  E.partials = nullptr; E.ticket = nullptr; E.result = nullptr; E.interior = nullptr; E.xr = nullptr; E.skip = nullptr;
  for (int d = 0; d < 4; d++) { E.X[d] = X[d]; E.dep[d] = 0; E.nbv[d] = 0; }
  E.Vh = Vh; E.parity = parity; E.ndir = ndir; E.seq = 1; E.wait = 0; E.err = nullptr; E.timeout_ns = 0;
  for (int i = 0; i < ndir; i++) {
    ExtDir& D = E.d[i];
    D.mu = halo[i].mu; D.depth = halo[i].depth; D.faceh = halo[i].faceh;
    D.nb = (2 * halo[i].depth < X[halo[i].mu]) ? 2 * halo[i].depth : X[halo[i].mu];
    D.ghost_f = (const double2*)recv + halo[i].off_f;
    D.ghost_b0 = (const double2*)recv + halo[i].off_b;
    D.ghost_b1 = D.ghost_b0 + (size_t)3 * halo[i].faceh;
    D.flag_f = D.flag_b = nullptr;
    E.dep[halo[i].mu] = halo[i].depth; E.nbv[halo[i].mu] = D.nb;
  }
  const int end = plan_regions(E);
  for (int t = 0; t < end; t++) { if (nl == 2) exterior_site<2, 0, false>(E, t); else exterior_site<1, 0, false>(E, t); }
  return 0;
}
// compress_links of layout.cu on the host: k_compress + k_sign_bwd site by site.  stats[6] = max deviation per class,
// min |s|^2, max |s|^2 (what the host decides on); u12_j sized [4][2][6 or 7][Vh] double2, sgn [2][Vh].
int emu_compress(int nl, const int* X, const double* links, double* u12_0, double* u12_1, unsigned short* sgn,
                 int mode0, int mode1, double* stats) {
  const int Vh = X[0] * X[1] * X[2] * X[3] / 2;
  double dev[2] = {0.0, 0.0}, smin[2] = {1e300, 1e300}, smax[2] = {0.0, 0.0};
  for (int par = 0; par < 2; par++)
    for (int idx = 0; idx < Vh; idx++) {
      unsigned int bits;
      if (nl == 2) bits = compress_site<2>((const double2*)links, (double2*)u12_0, (double2*)u12_1, Vh, par, idx, mode0, mode1, dev, smin, smax);
      else bits = compress_site<1>((const double2*)links, (double2*)u12_0, (double2*)u12_1, Vh, par, idx, mode0, 0, dev, smin, smax);
      if (sgn) sgn[(size_t)par * Vh + idx] = (unsigned short)bits;
    }
  if (sgn)
    for (int par = 0; par < 2; par++)      // all low bytes are final before any high byte is built (two launches on the GPU)
      for (int idx = 0; idx < Vh; idx++) {
        const size_t ds = (size_t)par * Vh + idx;
        unsigned int bits = sgn[ds] & 0xffu;
        bits = (nl == 2) ? sign_bwd_site<2>(sgn, Vh, par, idx, X, bits) : sign_bwd_site<1>(sgn, Vh, par, idx, X, bits);
        sgn[ds] = (unsigned short)bits;
      }
  for (int j = 0; j < 2; j++) { stats[j] = dev[j]; stats[2 + j] = smin[j]; stats[4 + j] = smax[j]; }
  return 0;
}
// idx of parity par -> local coordinates (site_from_idx): the map the tests check against the oracle's site order
void emu_site(int idx, int par, const int* X, int* out5) {
  const Site s = site_from_idx(idx, par, X);
  for (int d = 0; d < 4; d++) out5[d] = s.x[d];
  out5[4] = s.lex;
}
}
