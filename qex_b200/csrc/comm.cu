// Halo exchange over NVLink with NCCL send/recv -- the B200 counterpart of QEX's
// paired QMP messages (src/layout/qshifts.nim:23-131, src/layout/shifts.nim:67-94,
// 254-285).  Per split direction and sweep a rank sends
//   to its - neighbour : x on its first `depth` layers            (forward shifts)
//   to its + neighbour : U_k^+ x on its last h_k layers, per link  (backward shifts,
//                        the product is formed at the source like stagD.nim:359-361)
// All messages of a sweep go out in one ncclGroup on a dedicated stream so the
// interior kernel overlaps the transfer; the exterior kernels wait on an event.
#include "stencil.cuh"
#include <cstdlib>
#include <cstring>

namespace qb {

int pack_faces(const Gauge& g, int q, const double2* x, unsigned int seq, cudaStream_t st);  // dslash.cu

static int rank_of(const Geom& g, const int* rc) {
  int r = 0;
  for (int d = 0; d < 4; d++) r = r * g.G[d] + ((rc[d] + g.G[d]) % g.G[d]);
  return r;
}

static int halo_nl = 0;  // links per direction the buffers were laid out for

// Pure host arithmetic (no CUDA): which faces this rank exchanges and where they sit in the
// send/recv buffers.  Exposed as qexb200_plan_halo so the pairing can be tested without a GPU.
int plan_halo(const Geom& g, int nl, std::vector<HaloDir>& dirs, size_t* total) {
  dirs.clear();
  size_t off = 0;
  for (int mu = 0; mu < 4; mu++) {
    if (!((g.comm_mask >> mu) & 1)) continue;
    HaloDir h;
    h.mu = mu;
    h.faceh = g.V / g.X[mu] / 2;
    h.depth = (nl == 2) ? 3 : 1;
    if (g.X[mu] < h.depth) return fail("local extent %d in split direction %d is smaller than the hop depth %d", g.X[mu], mu, h.depth);
    h.n_f = (size_t)3 * h.depth * h.faceh;
    h.n_b = (size_t)3 * ((nl == 2) ? 4 : 1) * h.faceh;
    h.off_f = off; off += h.n_f;
    h.off_b = off; off += h.n_b;
    int rc[4] = {g.rc[0], g.rc[1], g.rc[2], g.rc[3]};
    rc[mu] = g.rc[mu] + 1; h.rank_plus = rank_of(g, rc);
    rc[mu] = g.rc[mu] - 1; h.rank_minus = rank_of(g, rc);
    dirs.push_back(h);
  }
  *total = off;
  return 0;
}

static int halo_layout(int nl) {
  Context& c = ctx();
  if (halo_nl == nl && (c.sendbuf || c.p2p || c.g.comm_mask == 0)) return 0;
  if (c.sendbuf) cudaFree(c.sendbuf);
  if (c.recvbuf) cudaFree(c.recvbuf);
  c.sendbuf = c.recvbuf = nullptr;
  size_t off = 0;
  QB_CHECK(plan_halo(c.g, nl, c.halo, &off));
  c.halo_elems = off;
  if (off && !c.p2p) {
    if (!c.nccl) return fail("multi-rank sweep needs either an NCCL communicator (qexb200_init with an id) or an attached peer window (qexb200_p2p_attach)");
    QB_CUDA(cudaStreamSynchronize(c.comm_stream));
    QB_CUDA(cudaMalloc(&c.sendbuf, sizeof(double2) * off));
    QB_CUDA(cudaMalloc(&c.recvbuf, sizeof(double2) * off));
  }
  if (c.p2p && off > c.halo_cap) return fail("halo of %zu elements exceeds the peer window (%zu)", off, c.halo_cap);
  halo_nl = nl;
  return 0;
}

int halo_setup() { return 0; }

int halo_free() {
  Context& c = ctx();
  if (c.sendbuf) cudaFree(c.sendbuf);
  if (c.recvbuf) cudaFree(c.recvbuf);
  c.sendbuf = c.recvbuf = nullptr;
  c.halo.clear();
  c.halo_elems = 0;
  halo_nl = 0;
  for (int i = 0; i < 2; i++)
    for (int p = 0; p < 2; p++) { if (c.d_order[i][p]) cudaFree(c.d_order[i][p]); c.d_order[i][p] = nullptr; }
  return 0;
}

// Diagnostic (qexb200_debug_role_times): on = 1 arms the timestamp ring of the multi-rank sweep and of the cross-rank
// sums; on = 0 copies it out (STAMP_RING launches x 8 words, then 8 words of xrank_sum) and disarms.
int role_times(int on, unsigned long long* out, int nwords) {
  Context& c = ctx();
  if (!c.p2p) return fail("role_times: needs the peer-memory transport");
  const size_t n = (size_t)STAMP_RING * 8 + 8;
  QB_CUDA(cudaStreamSynchronize(c.stream));
  if (on) {
    if (!c.d_stamps) QB_CUDA(cudaMalloc(&c.d_stamps, n * sizeof(unsigned long long)));
    std::vector<unsigned long long> h(n, 0ull);
    for (int i = 0; i < STAMP_RING; i++) h[(size_t)i * 8 + 0] = h[(size_t)i * 8 + 3] = ~0ull;
    QB_CUDA(cudaMemcpy(c.d_stamps, h.data(), n * sizeof(unsigned long long), cudaMemcpyHostToDevice));
    unsigned long long* xs = c.d_stamps + (size_t)STAMP_RING * 8;
    QB_CUDA(cudaMemcpy(&c.d_xr->stamps, &xs, sizeof xs, cudaMemcpyHostToDevice));
    return 0;
  }
  if (!c.d_stamps) return fail("role_times: not armed");
  if (out) {
    const size_t m = (size_t)nwords < n ? (size_t)nwords : n;
    QB_CUDA(cudaMemcpy(out, c.d_stamps, m * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  }
  unsigned long long* none = nullptr;
  QB_CUDA(cudaMemcpy(&c.d_xr->stamps, &none, sizeof none, cudaMemcpyHostToDevice));
  cudaFree(c.d_stamps);
  c.d_stamps = nullptr;
  return 0;
}

// ---- peer-memory window ---------------------------------------------------------------------
int p2p_alloc() {
  Context& c = ctx();
  if (c.g.nranks < 2 || c.win) return 0;
  std::vector<HaloDir> dirs;
  size_t cap = 0;
  bool naik_ok = true;   // size for the deeper (fat + Naik) halo when the local extents allow it
  for (int mu = 0; mu < 4; mu++)
    if (((c.g.comm_mask >> mu) & 1) && c.g.X[mu] < 3) naik_ok = false;
  QB_CHECK(plan_halo(c.g, naik_ok ? 2 : 1, dirs, &cap));
  c.halo_cap = cap;
  const size_t head = ((WIN_HDR + 2 * cap * sizeof(double2) + 255) / 256) * 256;
  // arena of link-shaped fields for the multi-rank link pipeline / force (QEXB200_ARENA_FIELDS matrix fields;
  // an nHYP smear holds 4 + 4 + 24, a HISQ smear 4 + 4 + 4 + 11)
  int units = 64;
  if (const char* ev = getenv("QEXB200_ARENA_FIELDS")) units = atoi(ev) < 0 ? 0 : atoi(ev);
  c.arena_unit = sizeof(double2) * 18 * (size_t)c.g.Vh;
  c.arena_off = head;
  c.win_bytes = head + (size_t)units * c.arena_unit;
  cudaError_t e = cudaMalloc(&c.win, c.win_bytes);
  if (e != cudaSuccess) return fail("p2p window of %zu bytes (%d arena fields, QEXB200_ARENA_FIELDS): %s", c.win_bytes, units, cudaGetErrorString(e));
  c.arena_units = units;
  c.arena_used.assign((size_t)units, 0);
  QB_CUDA(cudaMemset(c.win, 0, head));
  QB_CUDA(cudaDeviceSynchronize());
  return 0;
}

bool in_arena(const void* p) {
  Context& c = ctx();
  const char* q = static_cast<const char*>(p);
  return c.win && q >= c.win + c.arena_off && q < c.win + c.win_bytes;
}
// first fit over unit slots: deterministic, so ranks making the same calls get the same offsets
int links_alloc(int nfields, double2** out, bool* arena) {
  Context& c = ctx();
  *arena = false;
  if (c.g.nranks > 1 && c.win && c.arena_units > 0) {
    for (int i = 0; i + nfields <= c.arena_units; i++) {
      int j = 0;
      while (j < nfields && !c.arena_used[(size_t)(i + j)]) j++;
      if (j == nfields) {
        for (j = 0; j < nfields; j++) c.arena_used[(size_t)(i + j)] = 1;
        *out = reinterpret_cast<double2*>(c.win + c.arena_off + (size_t)i * c.arena_unit);
        *arena = true;
        return 0;
      }
      i += j;
    }
    return fail("the peer arena (%d matrix fields) is full: raise QEXB200_ARENA_FIELDS", c.arena_units);
  }
  QB_CUDA(cudaMalloc(out, sizeof(double2) * 18 * (size_t)c.g.Vh * nfields));
  return 0;
}
void links_free(double2* p, int nfields, bool arena) {
  Context& c = ctx();
  if (!p) return;
  if (!arena) { cudaFree(p); return; }
  if (!c.win) return;   // window already gone (finalize)
  const size_t i = (size_t)(reinterpret_cast<char*>(p) - (c.win + c.arena_off)) / c.arena_unit;
  for (int j = 0; j < nfields && i + j < c.arena_used.size(); j++) c.arena_used[i + j] = 0;
}
// rank that owns the sites displaced by nb = sum_d (disp_d + 1) 3^d rank-grid steps from this rank (periodic)
void plan_peers(const Geom& g, unsigned char peer_of[81]) {
  for (int nb = 0; nb < 81; nb++) {
    int t = nb, r = 0, rc[4];
    for (int d = 0; d < 4; d++) { const int disp = t % 3 - 1; t /= 3; rc[d] = ((g.rc[d] + disp) % g.G[d] + g.G[d]) % g.G[d]; }
    for (int d = 0; d < 4; d++) r = r * g.G[d] + rc[d];   // dimension 0 slowest (qlayout.nim:94-113)
    peer_of[nb] = (unsigned char)r;
  }
}
int peers_fill(Peers* pr) {
  Context& c = ctx();
  if (!c.p2p) return fail("this operation on a lattice split over ranks needs the peer-memory transport (qexb200_p2p_attach)");
  memset(pr, 0, sizeof *pr);
  for (int r = 0; r < c.g.nranks; r++) pr->base[r] = c.peer_win[r] + c.arena_off;
  pr->me = c.g.rank;
  plan_peers(c.g, pr->peer_of);
  return 0;
}
// Cross-rank barrier on the compute stream: one warp pushes a sequence word into every peer's LL slot and
// waits for everybody's (xrank_sum of zeros).  Kernel boundaries order it after this rank's earlier kernels.
__global__ void k_xbarrier(XRank* xr) { (void)xrank_sum(xr, 0.0); }
int xbarrier() {
  Context& c = ctx();
  if (c.g.nranks < 2) return 0;
  if (!c.p2p) return fail("cross-rank barrier needs the peer-memory transport (qexb200_p2p_attach)");
  k_xbarrier<<<1, 32, 0, c.stream>>>(c.d_xr);
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

int p2p_export(void* out64) {
  Context& c = ctx();
  if (c.g.nranks < 2) return fail("p2p_export: single-rank context has no peer window");
  QB_CHECK(p2p_alloc());
  cudaIpcMemHandle_t h;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t size");
  QB_CUDA(cudaIpcGetMemHandle(&h, c.win));
  memcpy(out64, &h, 64);
  return 0;
}

int p2p_attach(const void* handles) {
  Context& c = ctx();
  if (c.g.nranks < 2) return 0;
  if (c.g.nranks > XR_MAXR) return fail("p2p_attach: at most %d ranks (one box)", XR_MAXR);
  if (c.p2p) return fail("p2p_attach: already attached");
  QB_CHECK(p2p_alloc());
  QB_CUDA(cudaStreamSynchronize(c.stream));
  QB_CUDA(cudaStreamSynchronize(c.comm_stream));
  const char* hb = static_cast<const char*>(handles);
  for (int r = 0; r < c.g.nranks; r++) {
    if (r == c.g.rank) { c.peer_win[r] = c.win; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, hb + (size_t)64 * r, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      for (int q = 0; q < r; q++) if (q != c.g.rank && c.peer_win[q]) { cudaIpcCloseMemHandle(c.peer_win[q]); c.peer_win[q] = nullptr; }
      return fail("p2p_attach: cannot map the window of rank %d: %s", r, cudaGetErrorString(e));
    }
    c.peer_win[r] = static_cast<char*>(p);
  }
  XRank xr;
  memset(&xr, 0, sizeof xr);
  for (int r = 0; r < c.g.nranks; r++) xr.ll[r] = reinterpret_cast<unsigned long long*>(c.peer_win[r] + WIN_FLAGS);
  xr.nranks = c.g.nranks; xr.rank = c.g.rank; xr.seq = 0; xr.err = c.d_err;
  if (const char* ev = getenv("QEXB200_P2P_TIMEOUT_S")) {
    const double sec = atof(ev);
    if (sec > 0) c.p2p_timeout_ns = (unsigned long long)(sec * 1e9);
  }
  xr.timeout_ns = c.p2p_timeout_ns;
  if (const char* ev = getenv("QEXB200_PACK_BLOCKS")) { if (atoi(ev) > 0) c.pack_blocks = atoi(ev); }
  if (!c.d_xr) QB_CUDA(cudaMalloc(&c.d_xr, sizeof(XRank)));
  QB_CUDA(cudaMemcpy(c.d_xr, &xr, sizeof xr, cudaMemcpyHostToDevice));
  // buffers laid out for the NCCL path are no longer needed
  QB_CHECK(halo_free());
  c.halo_seq = 0;
  c.p2p = true;
  return 0;
}

int p2p_free() {
  Context& c = ctx();
  for (int r = 0; r < XR_MAXR; r++) {
    if (c.peer_win[r] && c.peer_win[r] != c.win) cudaIpcCloseMemHandle(c.peer_win[r]);
    c.peer_win[r] = nullptr;
  }
  if (c.win) cudaFree(c.win);
  if (c.d_xr) cudaFree(c.d_xr);
  if (c.d_stamps) cudaFree(c.d_stamps);
  c.d_stamps = nullptr;
  c.win = nullptr; c.d_xr = nullptr; c.p2p = false; c.win_bytes = 0; c.halo_cap = 0;
  c.arena_units = 0; c.arena_used.clear(); c.arena_off = 0;
  return 0;
}

// The exchange of one parity sweep.  Both paths pack on the comm stream, concurrently with the
// interior kernel on the compute stream:
//   peer-memory : k_pack stores straight into the neighbours' windows and raises their flags;
//                 the exterior kernel spins on its own flags (no host or NCCL involvement);
//   NCCL        : k_pack fills the local send buffer, one ncclGroup of send/recv follows.
// Lay out the halo for this link set and open exchange number halo_seq.
int halo_begin(const Gauge& g) {
  QB_CHECK(halo_layout(g.nlinks / 4));
  ctx().halo_seq++;
  return 0;
}

int halo_start(const Gauge& g, int parity, const double2* x) {
  Context& c = ctx();
  QB_CHECK(halo_begin(g));
  QB_CUDA(cudaEventRecord(c.ev_pack, c.stream));            // x is ready
  QB_CUDA(cudaStreamWaitEvent(c.comm_stream, c.ev_pack, 0));
  QB_CHECK(pack_faces(g, 1 - parity, x, c.halo_seq, c.comm_stream));
  if (!c.p2p) {
    QB_NCCL(nccl().GroupStart());
    // order matters when + and - neighbour are the same rank (rank grid of 2): the first
    // send to a peer pairs with that peer's first receive.
    for (const HaloDir& h : c.halo) {
      QB_NCCL(nccl().Send(c.sendbuf + h.off_f, h.n_f * 2, ncclDouble, h.rank_minus, c.nccl, c.comm_stream));
      QB_NCCL(nccl().Send(c.sendbuf + h.off_b, h.n_b * 2, ncclDouble, h.rank_plus, c.nccl, c.comm_stream));
      QB_NCCL(nccl().Recv(c.recvbuf + h.off_f, h.n_f * 2, ncclDouble, h.rank_plus, c.nccl, c.comm_stream));
      QB_NCCL(nccl().Recv(c.recvbuf + h.off_b, h.n_b * 2, ncclDouble, h.rank_minus, c.nccl, c.comm_stream));
    }
    QB_NCCL(nccl().GroupEnd());
  }
  QB_CUDA(cudaEventRecord(c.ev_comm, c.comm_stream));
  return 0;
}

int halo_wait() {
  Context& c = ctx();
  QB_CUDA(cudaStreamWaitEvent(c.stream, c.ev_comm, 0));
  return 0;
}

}  // namespace qb
