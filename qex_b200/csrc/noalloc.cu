// qexb200_norm2_noalloc: |x|^2 of a caller-owned device array without a single allocation.
//
// Who this is for.  Everything else in this library gets its working memory from cudaMalloc:
// qexb200_init builds a context (streams, reduction partials, pinned scalars), qexb200_field_create
// allocates the vectors the existing norm2 (qexb200_field_norm2 -> k_reduce in blas.cu) runs on.  Some
// users of a lattice code cannot do that at the moment they want a norm:
//   - a run that was sized to fill the GPU -- resident link sets of a 64^3 x 128 lattice, an arena handed to
//     another backend's memory pool, a job sharing the device with a process that owns the rest -- where a
//     late cudaMalloc can simply fail, in the middle of a trajectory;
//   - code inside a section where the device allocator's side effects are not acceptable: cudaMalloc /
//     cudaFree may synchronise the whole device, which stalls (or, with peer kernels spinning on flags,
//     can wedge) work in flight on other streams;
//   - monitoring or checking code that holds only raw device pointers from its own allocator and has no
//     qexb200 context at all (before qexb200_init, after qexb200_finalize, or in a process that never
//     calls them).
// For them the per-block partial sums and the completion counter of the reduction -- the scratch that
// k_reduce takes from the context -- live in the file-scope array below, which exists from the moment the
// library is loaded on a GPU until the process exits.
//
// No counterpart in the reference (QEX's norm2 is src/field/fieldET.nim:605-625 on host fields).
#include "common.cuh"

#include <mutex>

#include "../../include/qex_b200.h"

namespace {

constexpr int NA_BLOCK = 256;                  // threads per block
constexpr int NA_PER_THREAD = 16;              // double2 per thread
constexpr long long NA_TILE2 = (long long)NA_BLOCK * NA_PER_THREAD;   // double2 per block (64 KB of input)
constexpr int NA_MAX_BLOCKS = 16384;           // one partial each
constexpr int NA_CTRL = 32;                    // doubles reserved in front of the partials (256 B): [0] = ticket
static_assert(2 * NA_TILE2 * NA_MAX_BLOCKS == QEXB200_NOALLOC_MAX_REALS, "header and kernel disagree on the maximum");

// The scratch: 256 B of control words, then one partial per block.  128.25 KB per GPU, held until exit.
// In HOST code this name is only a handle for cudaGetSymbolAddress; device code may name it directly.
static __device__ __align__(256) double s_noalloc_scratch[NA_CTRL + NA_MAX_BLOCKS];

// Block b sums the fixed tile [b * NA_TILE2, (b+1) * NA_TILE2) of double2, so the summation order -- and the
// result, bit for bit -- depends on n only, not on the GPU's SM count.  The block that draws the last ticket adds
// the partials in index order.
__global__ void __launch_bounds__(NA_BLOCK) k_norm2_noalloc(const double* __restrict__ x, long long nreals,
                                                            double* scratch, double* out) {
  __shared__ double sm[NA_BLOCK / 32];
  __shared__ bool last;
  const double2* x2 = reinterpret_cast<const double2*>(x);
  const long long n2 = nreals >> 1;
  const long long base = (long long)blockIdx.x * NA_TILE2 + threadIdx.x;
  double2 v[NA_PER_THREAD];
#pragma unroll
  for (int k = 0; k < NA_PER_THREAD; k++) {
    const long long i = base + (long long)k * NA_BLOCK;
    v[k] = (i < n2) ? x2[i] : make_double2(0.0, 0.0);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < NA_PER_THREAD; k++) { s = fma(v[k].x, v[k].x, s); s = fma(v[k].y, v[k].y, s); }
  if ((nreals & 1) && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) { const double t = x[nreals - 1]; s = fma(t, t, s); }

#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) sm[w] = s;
  __syncthreads();
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
  double* partials = scratch + NA_CTRL;
  if (w == 0) {
    s = (lane < NA_BLOCK / 32) ? sm[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) {
      partials[blockIdx.x] = s;
      __threadfence();
      last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  s = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += NA_BLOCK) s += __ldcg(partials + i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  __syncthreads();
  if (lane == 0) sm[w] = s;
  __syncthreads();
  if (w == 0) {
    s = (lane < NA_BLOCK / 32) ? sm[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if (lane == 0) *out = s;
  }
}

// Device address of the scratch, per device, resolved once with that device current.
constexpr int NA_MAX_DEVICES = 64;
double* s_scratch_ptr[NA_MAX_DEVICES] = {};
std::mutex s_setup_mutex;

int scratch_for_current_device(double** ptr) {
  int dev = -1;
  QB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NA_MAX_DEVICES) return qb::fail("norm2_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(s_setup_mutex);
  if (!s_scratch_ptr[dev]) {
    double* p = nullptr;
    QB_CUDA(cudaGetSymbolAddress((void**)&p, s_noalloc_scratch));   // device `dev` is current here
    if (!p) return qb::fail("norm2_noalloc: cudaGetSymbolAddress returned a null scratch address on device %d", dev);
    s_scratch_ptr[dev] = p;
  }
  *ptr = s_scratch_ptr[dev];
  return 0;
}

int must_be_device_memory(const void* p, const char* what) {
  cudaPointerAttributes at;
  QB_CUDA(cudaPointerGetAttributes(&at, p));
  if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)
    return qb::fail("norm2_noalloc: %s must point to device memory (this is host memory)", what);
  return 0;
}

}  // namespace

extern "C" int qexb200_norm2_noalloc(const double* d_x, long long nreals, double* d_out, void* stream) {
  qb::clear_error();
  if (nreals < 0) return qb::fail("norm2_noalloc: nreals = %lld is negative", nreals);
  if (nreals > QEXB200_NOALLOC_MAX_REALS)
    return qb::fail("norm2_noalloc: nreals = %lld exceeds QEXB200_NOALLOC_MAX_REALS = %lld (the static scratch holds "
                    "%d partial sums of %lld reals each); split the array and add the pieces",
                    nreals, (long long)QEXB200_NOALLOC_MAX_REALS, NA_MAX_BLOCKS, 2 * NA_TILE2);
  if (!d_out) return qb::fail("norm2_noalloc: d_out is NULL");
  if (nreals > 0 && !d_x) return qb::fail("norm2_noalloc: d_x is NULL");
  if (((size_t)d_x & 15) != 0) return qb::fail("norm2_noalloc: d_x must be 16-byte aligned");
  if (((size_t)d_out & 7) != 0) return qb::fail("norm2_noalloc: d_out must be 8-byte aligned");
  QB_CHECK(must_be_device_memory(d_out, "d_out"));
  if (nreals > 0) QB_CHECK(must_be_device_memory(d_x, "d_x"));
  cudaStream_t st = (cudaStream_t)stream;
  if (nreals == 0) {   // empty subset: the sum is +0.0, whose bytes are all zero
    QB_CUDA(cudaMemsetAsync(d_out, 0, sizeof(double), st));
    return 0;
  }
  double* scratch = nullptr;
  QB_CHECK(scratch_for_current_device(&scratch));
  // The array keeps whatever the previous call (or a call that was cut short) left: the ticket must start at zero.
  QB_CUDA(cudaMemsetAsync(scratch, 0, NA_CTRL * sizeof(double), st));
  const long long n2 = nreals >> 1;
  const long long blocks = n2 > 0 ? (n2 + NA_TILE2 - 1) / NA_TILE2 : 1;
  k_norm2_noalloc<<<(unsigned int)blocks, NA_BLOCK, 0, st>>>(d_x, nreals, scratch, d_out);
  QB_CUDA(cudaGetLastError());
  return 0;
}
