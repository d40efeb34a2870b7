// Device helpers of the staggered stencil: site <-> coordinate maps on the
// checkerboarded local lattice, SU(3)-shaped complex mat-vec, streaming loads.
#pragma once
#include "common.cuh"

namespace qb {

struct Site {
  int x[4];
  int lex;
};

// idx in [0,Vh) of parity `par` -> local coordinates and lexicographic index
__device__ __forceinline__ Site site_from_idx(int idx, int par, const int* X) {
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

__device__ __forceinline__ int stride_of(int mu, const int* X) {
  return mu == 0 ? 1 : (mu == 1 ? X[0] : (mu == 2 ? X[0] * X[1] : X[0] * X[1] * X[2]));
}

// face index (other three coordinates, lower dimension fastest) >> 1
__device__ __forceinline__ int face_lex(const int* x, int mu, const int* X) {
  int l = 0;
#pragma unroll
  for (int d = 3; d >= 0; d--)
    if (d != mu) l = l * X[d] + x[d];
  return l;
}

// (layer coordinate xmu, face half index, parity) -> full coordinates
__device__ __forceinline__ Site site_from_face(int fidx, int mu, int xmu, int par, const int* X) {
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
__device__ __forceinline__ void load_link(double2 (&U)[9], const double2* __restrict__ p, int Vh) {
#pragma unroll
  for (int e = 0; e < 9; e++) U[e] = __ldcs(p + (size_t)e * Vh);
}
// conj(row0 x row1) of the two stored rows
__device__ __forceinline__ void cross_conj(double2 (&c)[3], const double2 (&U)[9]) {
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
__device__ __forceinline__ void load_link_k(double2 (&U)[9], const Links& L, int mu, int j, int par,
                                            int idx, int Vh, unsigned int neg) {
  if ((RC >> (2 + j)) & 1) {
    const double2* __restrict__ p = L.u12[j] + ((size_t)(mu * 2 + par) * 7) * Vh + idx;
#pragma unroll
    for (int e = 0; e < 6; e++) U[e] = __ldcs(p + (size_t)e * Vh);
    const double2 phi = __ldcs(p + (size_t)6 * Vh);
    double2 c[3];
    cross_conj(c, U);
#pragma unroll
    for (int e = 0; e < 3; e++) U[6 + e] = make_double2(phi.x * c[e].x - phi.y * c[e].y, phi.x * c[e].y + phi.y * c[e].x);
  } else if ((RC >> j) & 1) {
    const double2* __restrict__ p = L.u12[j] + ((size_t)(mu * 2 + par) * 6) * Vh + idx;
#pragma unroll
    for (int e = 0; e < 6; e++) U[e] = __ldcs(p + (size_t)e * Vh);
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
    for (int e = 0; e < 9; e++) U[e] = __ldcs(p + (size_t)e * Vh);
  }
}

// vectors are re-read by up to 16 neighbours: keep them cached
__device__ __forceinline__ void load_vec(double2 (&v)[3], const double2* __restrict__ p, int Vh) {
#pragma unroll
  for (int c = 0; c < 3; c++) v[c] = __ldg(p + (size_t)c * Vh);
}

// acc += U v
__device__ __forceinline__ void mv_add(double2 (&acc)[3], const double2 (&U)[9], const double2 (&v)[3]) {
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
__device__ __forceinline__ void mv_adj_sub(double2 (&acc)[3], const double2 (&U)[9], const double2 (&v)[3]) {
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

__device__ __forceinline__ const double2* link_ptr(const double2* u, int k, int par, int Vh, int idx) {
  return u + ((size_t)(k * 2 + par) * 9) * Vh + idx;
}

// ---- system-scope primitives for peer-memory signalling ---------------------------------
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Wait until *flag == want (written by a peer GPU with st.release.sys).  Bounded: after
// timeout_ns the mapped error word is set and the wait gives up, so a dead peer turns into an
// error status on the host instead of a hung GPU.
__device__ __forceinline__ void wait_flag(const unsigned int* flag, unsigned int want, int* err,
                                          unsigned long long timeout_ns) {
  if (ld_acquire_sys(flag) == want) return;
  const unsigned long long t0 = globaltimer_ns();
  while (ld_acquire_sys(flag) != want) {
    if (globaltimer_ns() - t0 > timeout_ns) {
      // diagnostics for the host: what was awaited, what was there, who waited
      err[1] = (int)want; err[2] = (int)ld_acquire_sys(flag); err[3] = (int)((blockIdx.x << 4) | (threadIdx.x & 15));
      err[0] = 1;
      __threadfence_system();
      break;
    }
  }
}

// Sum of one double over the ranks of the box, executed by ONE warp (all 32 lanes call it).
// Lane j < nranks pushes this rank's value into rank j's LL slot as two 8-byte words
// {seq | 32 bits of the double} (8-byte stores are single NVLink transactions, so a word whose
// seq matches carries valid data: no fence, one fabric traversal), then collects rank j's value
// from the local slot.  Every rank adds the same values in rank order: the result is
// bit-identical everywhere and run-to-run.
template <bool MAXOP = false>
__device__ __forceinline__ double xrank_sum(XRank* xr, double v) {
  const int lane = threadIdx.x & 31;
  const int n = xr->nranks, me = xr->rank;
  // lane 0 owns the counter; the others take its value (several sums may follow each other inside one kernel)
  unsigned int seq = 0;
  if (lane == 0) seq = *(volatile unsigned int*)&xr->seq + 1u;
  seq = __shfl_sync(0xffffffffu, seq, 0);
  const int b = seq % XR_NB;
  double got = 0.0;
  unsigned long long* const stamps = xr->stamps;
  const unsigned long long t_in = stamps ? globaltimer_ns() : 0ull;
  if (lane < n) {
    const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
    const unsigned long long hi = (unsigned long long)seq << 32;
    unsigned long long* dst = xr->ll[lane] + ((size_t)(b * XR_MAXR + me) * 2);
    st_relaxed_sys64(dst, hi | (bits & 0xffffffffull));
    st_relaxed_sys64(dst + 1, hi | (bits >> 32));
    const unsigned long long* src = xr->ll[me] + ((size_t)(b * XR_MAXR + lane) * 2);
    unsigned long long w0 = ld_relaxed_sys64(src), w1 = ld_relaxed_sys64(src + 1);
    if ((w0 >> 32) != seq || (w1 >> 32) != seq) {
      const unsigned long long t0 = globaltimer_ns();
      for (;;) {
        w0 = ld_relaxed_sys64(src); w1 = ld_relaxed_sys64(src + 1);
        if ((w0 >> 32) == seq && (w1 >> 32) == seq) break;
        if (globaltimer_ns() - t0 > xr->timeout_ns) { *xr->err = 2; __threadfence_system(); break; }
      }
    }
    got = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
  }
  double s = MAXOP ? -1.797693134862315e308 : 0.0;
  for (int j = 0; j < n; j++) {
    const double gj = __shfl_sync(0xffffffffu, got, j);
    s = MAXOP ? fmax(s, gj) : s + gj;
  }
  __syncwarp();
  if (lane == 0) {
    *(volatile unsigned int*)&xr->seq = seq;
    if (stamps) {
      const unsigned long long dt = globaltimer_ns() - t_in;
      stamps[0] += dt; stamps[1] += 1ull;
      if (dt > stamps[2]) stamps[2] = dt;
    }
  }
  return s;
}

// Deterministic grid reduction: per-block partial -> last block sums in fixed order; *addend
// (a partial sum left by an earlier kernel) is added next; with xr != nullptr the last block
// also sums over the ranks (xrank_sum).  Returns true on the one thread that stored *result
// (it runs after every block has finished its loads, so it may update loop-control scalars).
template <int BLOCK>
__device__ __forceinline__ bool grid_reduce(double v, double* partials, unsigned int* ticket,
                                            double* result, XRank* xr = nullptr,
                                            const double* addend = nullptr) {
  __shared__ double sm[BLOCK / 32];
  __shared__ bool last;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) sm[w] = v;
  __syncthreads();
  if (w == 0) {
    v = (lane < BLOCK / 32) ? sm[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) {
      partials[blockIdx.x] = v;
      __threadfence();
      unsigned int t = atomicAdd(ticket, 1u);
      last = (t == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (last) {
    __threadfence();
    double s = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += BLOCK) s += __ldcg(partials + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    __syncthreads();
    if (lane == 0) sm[w] = s;
    __syncthreads();
    if (w == 0) {
      s = (lane < BLOCK / 32) ? sm[lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
      s = __shfl_sync(0xffffffffu, s, 0);
      if (addend) s += *addend;
      if (xr) s = xrank_sum(xr, s);
      if (lane == 0) {
        *result = s;
        *ticket = 0u;
        return true;
      }
    }
  }
  return false;
}

}  // namespace qb
