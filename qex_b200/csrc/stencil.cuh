// Device helpers of the staggered stencil: site <-> coordinate maps on the
// checkerboarded local lattice, SU(3)-shaped complex mat-vec, streaming loads.
#pragma once
#include "common.cuh"

namespace qb {

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
