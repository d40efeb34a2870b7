// Field BLAS-1 and reductions on device half/full fields (what QEX's expression
// templates do on a subset: src/field/fieldET.nim:547-599 assignments,
// :605-625 norm2, :704-724 redot).  Pure HBM streaming: 128-bit loads, grid sized
// to a few waves of 148 SMs, deterministic two-level reductions.
#include "stencil.cuh"

namespace qb {

constexpr int BB = 256;

static inline int blas_grid(size_t n) {
  size_t g = (n + BB - 1) / BB;
  const size_t cap = 148 * 8;
  return (int)(g < cap ? (g ? g : 1) : cap);
}

enum { OP_ZERO, OP_COPY, OP_SCALE, OP_AXPY, OP_XPAY, OP_XMY, OP_AXPBY };

template <int OP>
__global__ void __launch_bounds__(BB) k_blas(double2* y, const double2* x,
                                              const double2* z, double a, double b, size_t n) {
  for (size_t i = (size_t)blockIdx.x * BB + threadIdx.x; i < n; i += (size_t)gridDim.x * BB) {
    double2 o;
    if (OP == OP_ZERO) o = make_double2(0, 0);
    if (OP == OP_COPY) o = x[i];
    if (OP == OP_SCALE) { o = y[i]; o.x *= a; o.y *= a; }
    if (OP == OP_AXPY) { o = y[i]; const double2 v = x[i]; o.x = fma(a, v.x, o.x); o.y = fma(a, v.y, o.y); }
    if (OP == OP_XPAY) { o = y[i]; const double2 v = x[i]; o.x = fma(a, o.x, v.x); o.y = fma(a, o.y, v.y); }
    if (OP == OP_XMY) { const double2 v = x[i], w = z[i]; o.x = v.x - w.x; o.y = v.y - w.y; }
    if (OP == OP_AXPBY) { o = y[i]; const double2 v = x[i]; o.x = a * v.x + b * o.x; o.y = a * v.y + b * o.y; }
    y[i] = o;
  }
}

#define LAUNCH_BLAS(OP, y, x, z, a, b, n)                                            \
  do {                                                                               \
    k_blas<OP><<<blas_grid(n), BB, 0, ctx().stream>>>(y, x, z, a, b, n);             \
    count_launch();                                                                  \
    QB_CUDA(cudaGetLastError());                                                     \
  } while (0)

int blas_zero(double2* x, size_t n) { LAUNCH_BLAS(OP_ZERO, x, nullptr, nullptr, 0, 0, n); return 0; }
int blas_copy(double2* y, const double2* x, size_t n) { LAUNCH_BLAS(OP_COPY, y, x, nullptr, 0, 0, n); return 0; }
int blas_scale(double2* x, double s, size_t n) { LAUNCH_BLAS(OP_SCALE, x, nullptr, nullptr, s, 0, n); return 0; }
int blas_axpy(double2* y, double a, const double2* x, size_t n) { LAUNCH_BLAS(OP_AXPY, y, x, nullptr, a, 0, n); return 0; }
int blas_xpay(double2* y, const double2* x, double a, size_t n) { LAUNCH_BLAS(OP_XPAY, y, x, nullptr, a, 0, n); return 0; }
int blas_xmy(double2* z, const double2* x, const double2* y, size_t n) { LAUNCH_BLAS(OP_XMY, z, x, y, 0, 0, n); return 0; }
int blas_axpby(double2* y, double a, const double2* x, double b, size_t n) { LAUNCH_BLAS(OP_AXPBY, y, x, nullptr, a, b, n); return 0; }

template <bool TWO>
__global__ void __launch_bounds__(BB) k_reduce(const double2* __restrict__ x, const double2* __restrict__ y,
                                                size_t n, double* partials, unsigned int* ticket, double* out,
                                                XRank* xr) {
  double s = 0.0;
  for (size_t i = (size_t)blockIdx.x * BB + threadIdx.x; i < n; i += (size_t)gridDim.x * BB) {
    const double2 v = x[i];
    if (TWO) { const double2 w = y[i]; s = fma(v.x, w.x, s); s = fma(v.y, w.y, s); }
    else { s = fma(v.x, v.x, s); s = fma(v.y, v.y, s); }
  }
  grid_reduce<BB>(s, partials, ticket, out, xr);
}

int blas_norm2(const double2* x, size_t n, double* d_out) {
  Context& c = ctx();
  k_reduce<false><<<blas_grid(n), BB, 0, c.stream>>>(x, nullptr, n, c.d_partials, c.d_ticket, d_out, xrank());
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}
int blas_redot(const double2* x, const double2* y, size_t n, double* d_out) {
  Context& c = ctx();
  k_reduce<true><<<blas_grid(n), BB, 0, c.stream>>>(x, y, n, c.d_partials, c.d_ticket, d_out, xrank());
  count_launch();
  QB_CUDA(cudaGetLastError());
  return 0;
}

// Sum over ranks of scalars produced by blas_norm2 / blas_redot / the fused CG kernels.  On the
// peer-memory path those kernels already summed over the ranks (xrank_sum in their last block).
int allreduce_scalars(double* d_ptr, int n) {
  Context& c = ctx();
  if (c.g.nranks > 1 && !c.p2p) QB_NCCL(nccl().AllReduce(d_ptr, d_ptr, n, ncclDouble, ncclSum, c.nccl, c.stream));
  return 0;
}

int fetch_scalars(const double* d_ptr, double* h_out, int n) {
  Context& c = ctx();
  QB_CUDA(cudaMemcpyAsync(c.h_scal, d_ptr, sizeof(double) * n, cudaMemcpyDeviceToHost, c.stream));
  QB_CUDA(cudaStreamSynchronize(c.stream));
  for (int i = 0; i < n; i++) h_out[i] = c.h_scal[i];
  return check_device_error();
}

int check_device_error() {
  Context& c = ctx();
  if (c.h_err && *(volatile int*)c.h_err != 0) {
    const int e = *(volatile int*)c.h_err;
    return fail("a device-side wait on a peer rank timed out (%s; awaited seq %d, found %d, waiter %d of exchange %u); the result is invalid",
                e == 1 ? "halo flag" : "cross-rank sum", c.h_err[1], c.h_err[2], c.h_err[3], c.halo_seq);
  }
  return 0;
}

}  // namespace qb
