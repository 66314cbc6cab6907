// Vector kernels of the PCG loop and the scalar updates that close each reduction (shared by the single-GPU
// solver, pcg.cu, and the row-partitioned one, dist.cu).
#pragma once
#include "csr.h"

namespace b200spai {

// ||b||^2 is known: tolerances, iteration 0 of the history, zero / NaN right-hand sides.
__device__ __forceinline__ void pcg_final_init(PcgScalars* sc, double total, double* hist) {
  const double bn = sqrt(total);
  sc->bnorm = bn;
  sc->atol = fmax(0.0, sc->rtol * bn);
  sc->rr = total;
  sc->rho = sc->rho_prev = sc->pq = sc->alpha = sc->beta = 0.0;
  sc->iter = 0;
  sc->breakdown = 0;
  const int zero_rhs = (bn == 0.0) || !(bn == bn);
  sc->done = zero_rhs;
  sc->converged = (bn == 0.0);
  if (!(bn == bn)) sc->breakdown = 1;
  hist[0] = bn;
  // graph driver: the WHILE node that follows must not be entered for b == 0 (every body kernel would return
  // at its `done` test and nobody would ever clear the loop condition)
  if (sc->use_cond) cudaGraphSetConditional((cudaGraphConditionalHandle)sc->cond_handle, zero_rhs ? 0u : 1u);
}

// ||r||^2 of the updated residual is known: iteration count, history, convergence test, loop condition.
__device__ __forceinline__ void pcg_final_rr(PcgScalars* sc, double total, double* hist) {
  if (sc->breakdown || !(total == total)) {
    sc->breakdown = 1;
    sc->done = 1;
  } else {
    sc->rr = total;
    sc->rho_prev = sc->rho;
    const int it = sc->iter + 1;
    sc->iter = it;
    const double nrm = sqrt(total);
    hist[it] = nrm;
    if (nrm < sc->atol) {
      sc->done = 1;
      sc->converged = 1;
    } else if (it >= sc->max_iter) {
      sc->done = 1;
    }
  }
  if (sc->use_cond) cudaGraphSetConditional((cudaGraphConditionalHandle)sc->cond_handle, sc->done ? 0u : 1u);
}

static constexpr int VTHREADS = 256;
static constexpr int MAX_PARTIALS = 8192;

static inline int vec_grid(int n) {
  long long g = ((long long)n + VTHREADS * 4 - 1) / (VTHREADS * 4);
  const long long cap = (long long)sm_count() * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

static __global__ void __launch_bounds__(VTHREADS) pcg_init_kernel(const double* __restrict__ b, double* __restrict__ x,
                                                            double* __restrict__ r, int n, PcgScalars* sc,
                                                            double* partials, unsigned* ticket, double* hist) {
  __shared__ double red[34];
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double v = b[i];
    r[i] = v;
    x[i] = 0.0;
    acc += v * v;
  }
  double cta = block_sum(acc, red);
  double total;
  if (grid_sum_last(cta, partials, ticket, red, total) && threadIdx.x == 0) {
    if (sc->dist) sc->local_total = total;  // row-partitioned solve: the all-reduce kernel that follows finishes the job
    else pcg_final_init(sc, total, hist);
  }
}

// z = r (d == nullptr) or z = r / d ; rho = r.z
static __global__ void __launch_bounds__(VTHREADS) pcg_jacobi_kernel(const double* __restrict__ r, const double* __restrict__ d,
                                                              double* __restrict__ z, int n, PcgScalars* sc,
                                                              double* partials, unsigned* ticket) {
  __shared__ double red[34];
  if (sc->done) return;
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double rv = r[i];
    const double zv = d ? rv / d[i] : rv;
    z[i] = zv;
    acc += rv * zv;
  }
  double cta = block_sum(acc, red);
  double total;
  if (grid_sum_last(cta, partials, ticket, red, total) && threadIdx.x == 0) pcg_final_scalar(sc, FINAL_RHO, total);
}

// The two vector kernels move 16 bytes per thread and access (double2) with two independent pairs in flight.
// lo_halo / hi_halo (row-partitioned solve): the same update is also applied to the halo entries [-lo_halo, 0) and
// [n, n + hi_halo) of the extended vectors, whose z arrived with the rho all-reduce -- identical roundings, so a halo
// entry carries the bits its owner computes and p needs no exchange of its own.
static __global__ void __launch_bounds__(VTHREADS) pcg_update_p_kernel(const double* __restrict__ z, double* __restrict__ p,
                                                                int n, const PcgScalars* sc, int lo_halo = 0,
                                                                int hi_halo = 0) {
  if (sc->done) return;
  const bool first = (sc->iter == 0);
  const double beta = sc->beta;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n2 = n >> 1;
  const double2* z2 = reinterpret_cast<const double2*>(z);
  double2* p2 = reinterpret_cast<double2*>(p);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    const double2 zv = z2[i];
    // scipy: p *= beta; p += z  (two roundings, reproduced: no fma contraction)
    double2 pv = zv;
    if (!first) {
      pv = p2[i];
      pv.x = __dadd_rn(__dmul_rn(pv.x, beta), zv.x);
      pv.y = __dadd_rn(__dmul_rn(pv.y, beta), zv.y);
    }
    p2[i] = pv;
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const double zv = z[n - 1];
    p[n - 1] = first ? zv : __dadd_rn(__dmul_rn(p[n - 1], beta), zv);
  }
  for (long long h = (long long)blockIdx.x * blockDim.x + threadIdx.x; h < (long long)lo_halo + hi_halo; h += stride) {
    const long long i = h < lo_halo ? h - lo_halo : n + (h - lo_halo);
    const double zv = z[i];
    p[i] = first ? zv : __dadd_rn(__dmul_rn(p[i], beta), zv);
  }
}

static __global__ void __launch_bounds__(VTHREADS) pcg_update_xr_kernel(double* __restrict__ x, double* __restrict__ r,
                                                                 const double* __restrict__ p, const double* __restrict__ q,
                                                                 int n, PcgScalars* sc, double* partials, unsigned* ticket,
                                                                 double* hist, int lo_halo = 0, int hi_halo = 0) {
  __shared__ double red[34];
  if (sc->done) {  // never inside the loop of a healthy solve; keeps a WHILE body from spinning if it happens
    if (sc->use_cond && blockIdx.x == 0 && threadIdx.x == 0)
      cudaGraphSetConditional((cudaGraphConditionalHandle)sc->cond_handle, 0u);
    return;
  }
  const double alpha = sc->alpha;
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n2 = n >> 1;
  double2* x2 = reinterpret_cast<double2*>(x);
  double2* r2 = reinterpret_cast<double2*>(r);
  const double2* p2 = reinterpret_cast<const double2*>(p);
  const double2* q2 = reinterpret_cast<const double2*>(q);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    // scipy: x += alpha*p ; r -= alpha*q  (product rounded, then add)
    double2 xv = x2[i], rv = r2[i];
    const double2 pv = p2[i], qv = q2[i];
    xv.x = __dadd_rn(xv.x, __dmul_rn(alpha, pv.x));
    xv.y = __dadd_rn(xv.y, __dmul_rn(alpha, pv.y));
    rv.x = __dsub_rn(rv.x, __dmul_rn(alpha, qv.x));
    rv.y = __dsub_rn(rv.y, __dmul_rn(alpha, qv.y));
    x2[i] = xv;
    r2[i] = rv;
    acc += rv.x * rv.x;
    acc += rv.y * rv.y;
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const int i = n - 1;
    x[i] = __dadd_rn(x[i], __dmul_rn(alpha, p[i]));
    const double rv = __dsub_rn(r[i], __dmul_rn(alpha, q[i]));
    r[i] = rv;
    acc += rv * rv;
  }
  // halo entries of r (row-partitioned solve; q's halo arrived with the p.q all-reduce): not part of ||r||^2
  for (long long h = (long long)blockIdx.x * blockDim.x + threadIdx.x; h < (long long)lo_halo + hi_halo; h += stride) {
    const long long i = h < lo_halo ? h - lo_halo : n + (h - lo_halo);
    r[i] = __dsub_rn(r[i], __dmul_rn(alpha, q[i]));
  }
  double cta = block_sum(acc, red);
  double total;
  if (grid_sum_last(cta, partials, ticket, red, total) && threadIdx.x == 0) {
    if (sc->dist) {  // all-reduced together with the next rho (one exchange instead of two, same arithmetic)
      sc->local_rr = total;
      sc->have_rr = 1;
    } else {
      pcg_final_rr(sc, total, hist);
    }
  }
}

}  // namespace b200spai
