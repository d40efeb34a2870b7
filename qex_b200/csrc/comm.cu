// Halo exchange over NVLink with NCCL send/recv -- the B200 counterpart of QEX's
// paired QMP messages (src/layout/qshifts.nim:23-131, src/layout/shifts.nim:67-94,
// 254-285).  Per split direction and sweep a rank sends
//   to its - neighbour : x on its first `depth` layers            (forward shifts)
//   to its + neighbour : U_k^+ x on its last h_k layers, per link  (backward shifts,
//                        the product is formed at the source like stagD.nim:359-361)
// All messages of a sweep go out in one ncclGroup on a dedicated stream so the
// interior kernel overlaps the transfer; the exterior kernels wait on an event.
#include "common.cuh"

namespace qb {

int pack_faces(const Gauge& g, int q, const double2* x);  // dslash.cu

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
  if (halo_nl == nl && (c.sendbuf || c.g.comm_mask == 0)) return 0;
  QB_CHECK(halo_free());
  size_t off = 0;
  QB_CHECK(plan_halo(c.g, nl, c.halo, &off));
  c.halo_elems = off;
  if (off) {
    QB_CUDA(cudaMalloc(&c.sendbuf, sizeof(double2) * off));
    QB_CUDA(cudaMalloc(&c.recvbuf, sizeof(double2) * off));
  }
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
  return 0;
}

int halo_start(const Gauge& g, int parity, const double2* x) {
  Context& c = ctx();
  QB_CHECK(halo_layout(g.nlinks / 4));
  QB_CHECK(pack_faces(g, 1 - parity, x));
  QB_CUDA(cudaEventRecord(c.ev_pack, c.stream));
  QB_CUDA(cudaStreamWaitEvent(c.comm_stream, c.ev_pack, 0));
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
  QB_CUDA(cudaEventRecord(c.ev_comm, c.comm_stream));
  return 0;
}

int halo_wait() {
  Context& c = ctx();
  QB_CUDA(cudaStreamWaitEvent(c.stream, c.ev_comm, 0));
  return 0;
}

}  // namespace qb
