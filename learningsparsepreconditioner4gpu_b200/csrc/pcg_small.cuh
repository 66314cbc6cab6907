// Whole PCG solve of a SMALL system in one cooperative kernel (BASELINE config C2: batches of 0.4 k - 32 k node
// meshes).  The 5-kernel iteration of pcg.cu is latency-bound there (~5 us per graph node, 25 us per iteration for a
// system whose SpMV takes ~1 us); here an iteration is 3 grid-wide barriers (2 without G):
//
//   P1   [x += alpha p ; r' = r - alpha q ; ||r'||^2]  and  t = G^T r'   -- r' of a NEIGHBOUR row is recomputed on the fly
//        from (r, q) with the same two roundings its owner uses, so the residual update needs no barrier of its own
//   --- barrier: ||r'||^2 reduced -> iteration count, history, convergence / max_iter test (scipy cg order) ---
//   P2   z = G t + eps r' [/ d] ; rho = r'.z
//   --- barrier: rho reduced -> beta ---
//   P3   p' = z + beta p  and  q = A p'  (p' of a neighbour column recomputed on the fly from (z, p)) ; p'.q
//   --- barrier: p.q reduced -> alpha (breakdown if <= 0) ---
//
// none / diagonal: P1 also forms z = r' [/ d] and r'.z, so one barrier serves both reductions.
// Rows are summed sequentially with a rounded product and a rounded add (scipy csr_matvec order, like the streaming
// kernel); dot products: fixed per-thread row order -> fixed block tree -> per-CTA partials summed in index order by
// every CTA redundantly, so all CTAs hold identical scalars and take identical branches without a broadcast.
// r and p are double-buffered (a CTA may still gather the old vector while its neighbour writes the new one).
#pragma once
#include <cooperative_groups.h>

#include "pcg_kernels.cuh"

namespace b200spai {

struct SmallArgs {
  CsrView A, G, GT;
  const double* b;
  double* x;
  double *r0, *r1, *p0, *p1, *z, *q, *t;
  const double* d;  // diagonal of A (diagonal / scaled SPAI) or nullptr
  double eps;
  int precond;
  PcgScalars* sc;
  double* partials;  // [3][gridDim.x] rotating reduction slots ([.][2 * gridDim.x] used by the two-value reduction)
  double* hist;
};

constexpr int SMALL_THREADS = 256;
constexpr int SMALL_MAX_CTAS = 1024;

// sum of `mine` over the grid, identical bits in every thread of every CTA; slot rotates per call
__device__ __forceinline__ double small_grid_sum(cooperative_groups::grid_group& grid, double mine, double* slot, double* red) {
  const double cta = block_sum(mine, red);
  if (threadIdx.x == 0) slot[blockIdx.x] = cta;
  grid.sync();
  // warp 0: lane-strided partial sums in index order, then a fixed shuffle tree
  if (threadIdx.x < 32) {
    double acc = 0.0;
    for (unsigned i = threadIdx.x; i < gridDim.x; i += 32) acc += __ldcg(slot + i);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) red[33] = acc;
  }
  __syncthreads();
  const double total = red[33];
  __syncthreads();
  return total;
}

// One row of M times a vector whose entry c is produced by `entry(c)`.  The row is walked in batches of 8 entries: the
// 8 column indices / values and then the 8 gathered vector entries are loaded before any of them is used, so a row costs
// ~2 L2 round trips instead of one per entry (the vectors change every iteration and are written by other SMs, so they
// are read with ld.cg from L2).  The sum itself stays sequential: rounded product, rounded add, in entry order.
template <class Entry>
__device__ __forceinline__ double small_row_gen(const CsrView& M, int row, Entry entry) {
  double sum = 0.0;
  const int e = M.rowptr[row + 1];
  for (int k = M.rowptr[row]; k < e; k += 8) {
    int c[8];
    double v[8], u[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const bool in = k + j < e;
      c[j] = in ? M.col[k + j] : 0;
      v[j] = in ? M.val[k + j] : 0.0;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) u[j] = (k + j < e) ? entry(c[j]) : 0.0;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (k + j < e) sum = __dadd_rn(sum, __dmul_rn(v[j], u[j]));
  }
  return sum;
}

__device__ __forceinline__ double small_row(const CsrView& M, int row, const double* __restrict__ x) {
  return small_row_gen(M, row, [x](int c) { return __ldcg(x + c); });
}

// row of M times the vector u[c] = a[c] - alpha * b[c] formed on the fly (r' = r - alpha q)
__device__ __forceinline__ double small_row_sub(const CsrView& M, int row, const double* __restrict__ a,
                                                const double* __restrict__ b, double alpha) {
  return small_row_gen(M, row, [a, b, alpha](int c) { return __dsub_rn(__ldcg(a + c), __dmul_rn(alpha, __ldcg(b + c))); });
}

// row of M times the vector u[c] = p[c] * beta + z[c] formed on the fly (p' = z + beta p; first iteration p' = z)
__device__ __forceinline__ double small_row_axpy(const CsrView& M, int row, const double* __restrict__ z,
                                                 const double* __restrict__ p, double beta, bool first) {
  if (first) return small_row_gen(M, row, [z](int c) { return __ldcg(z + c); });
  return small_row_gen(M, row, [z, p, beta](int c) { return __dadd_rn(__dmul_rn(__ldcg(p + c), beta), __ldcg(z + c)); });
}

__global__ void __launch_bounds__(SMALL_THREADS) pcg_small_kernel(SmallArgs a) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ double red[34];
  const int n = a.A.n_rows;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int gstride = gridDim.x * blockDim.x;
  const bool spai = a.precond >= B200SPAI_PRECOND_EXT_SPAI;
  const bool lead = (blockIdx.x == 0 && threadIdx.x == 0);
  PcgScalars* sc = a.sc;
  const double rtol = sc->rtol;
  const int max_iter = sc->max_iter;
  double* slot[3] = {a.partials, a.partials + 2 * gridDim.x, a.partials + 4 * gridDim.x};
  int ns = 0;

  // ---- init: r = b, x = 0, ||b|| ---------------------------------------------------------------------
  double acc = 0.0;
  for (int i = gtid; i < n; i += gstride) {
    const double v = a.b[i];
    a.r0[i] = v;
    a.x[i] = 0.0;
    acc += v * v;
  }
  const double bb = small_grid_sum(grid, acc, slot[ns], red);
  ns = (ns + 1) % 3;
  const double bnorm = sqrt(bb);
  const double atol = fmax(0.0, rtol * bnorm);
  double rr = bb, rho = 0.0, rho_prev = 0.0, pq = 0.0, alpha = 0.0, beta = 0.0;
  int iter = 0, done = 0, converged = 0, breakdown = 0;
  if (lead) a.hist[0] = bnorm;
  if (bnorm == 0.0) {
    done = 1;
    converged = 1;
  } else if (!(bnorm == bnorm)) {
    done = 1;
    breakdown = 1;
  }
  double *r = a.r0, *rn = a.r1, *p = a.p0, *pn = a.p1;
  bool pending = false;  // an (alpha, p, q) update of x and r is waiting to be applied in P1

  while (!done) {
    // ---- P1 -------------------------------------------------------------------------------------------
    double rr_part = 0.0, rz_part = 0.0;
    for (int i = gtid; i < n; i += gstride) {
      double rv = r[i];
      if (pending) {
        a.x[i] = __dadd_rn(a.x[i], __dmul_rn(alpha, p[i]));
        rv = __dsub_rn(rv, __dmul_rn(alpha, a.q[i]));
        rn[i] = rv;
        rr_part += rv * rv;
      }
      if (spai) {
        double tv = pending ? small_row_sub(a.GT, i, r, a.q, alpha) : small_row(a.GT, i, r);
        if (a.precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) tv = tv / a.d[i];
        a.t[i] = tv;
      } else {
        const double zv = (a.precond == B200SPAI_PRECOND_DIAGONAL) ? rv / a.d[i] : rv;
        a.z[i] = zv;
        rz_part += rv * zv;
      }
    }
    if (pending) {
      double* tmp = r;
      r = rn;
      rn = tmp;
    }
    if (pending || !spai) {  // one barrier for ||r||^2 (and r.z without G)
      // two values through one slot: [0, grid) and [grid, 2 grid)
      const double c0 = block_sum(rr_part, red);
      __syncthreads();
      const double c1 = block_sum(rz_part, red);
      if (threadIdx.x == 0) {
        slot[ns][blockIdx.x] = c0;
        slot[ns][gridDim.x + blockIdx.x] = c1;
      }
      grid.sync();
      if (threadIdx.x < 32) {
        double s0 = 0.0, s1 = 0.0;
        for (unsigned i = threadIdx.x; i < gridDim.x; i += 32) {
          s0 += __ldcg(slot[ns] + i);
          s1 += __ldcg(slot[ns] + gridDim.x + i);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          s0 += __shfl_xor_sync(0xffffffffu, s0, o);
          s1 += __shfl_xor_sync(0xffffffffu, s1, o);
        }
        if (threadIdx.x == 0) {
          red[32] = s0;
          red[33] = s1;
        }
      }
      __syncthreads();
      const double rr_new = red[32], rz = red[33];
      __syncthreads();
      ns = (ns + 1) % 3;
      if (pending) {  // the update is complete: count it, test it (scipy: callback, then the test at the loop top)
        if (!(rr_new == rr_new)) {
          breakdown = 1;
          done = 1;
        } else {
          rr = rr_new;
          rho_prev = rho;
          iter += 1;
          const double nrm = sqrt(rr_new);
          if (lead) a.hist[iter] = nrm;
          if (nrm < atol) {
            done = 1;
            converged = 1;
          } else if (iter >= max_iter) {
            done = 1;
          }
        }
      }
      if (!spai) rho = rz;
    } else {
      grid.sync();  // first iteration with G: only t has to be complete
    }
    if (done) break;
    // ---- P2 (SPAI): z = G t + eps r [/ d], rho = r.z ------------------------------------------------------
    if (spai) {
      double part = 0.0;
      for (int i = gtid; i < n; i += gstride) {
        const double rv = r[i];
        double e = a.eps * rv;
        if (a.precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) e = e / a.d[i];
        const double zv = small_row(a.G, i, a.t) + e;
        a.z[i] = zv;
        part += rv * zv;
      }
      rho = small_grid_sum(grid, part, slot[ns], red);
      ns = (ns + 1) % 3;
    }
    const bool first = (iter == 0);
    beta = first ? 0.0 : rho / rho_prev;
    // ---- P3: p' = z + beta p, q = A p', p'.q --------------------------------------------------------------
    double part = 0.0;
    for (int i = gtid; i < n; i += gstride) {
      const double zv = a.z[i];
      const double pv = first ? zv : __dadd_rn(__dmul_rn(p[i], beta), zv);
      pn[i] = pv;
      const double qv = small_row_axpy(a.A, i, a.z, p, beta, first);
      a.q[i] = qv;
      part += pv * qv;
    }
    pq = small_grid_sum(grid, part, slot[ns], red);
    ns = (ns + 1) % 3;
    {
      double* tmp = p;
      p = pn;
      pn = tmp;
    }
    if (!(pq > 0.0) || !(rho == rho)) {  // p.q <= 0 or NaN: A or M not SPD
      breakdown = 1;
      alpha = 0.0;
      done = 1;
      break;
    }
    alpha = rho / pq;
    pending = true;
  }

  if (lead) {
    sc->bnorm = bnorm;
    sc->atol = atol;
    sc->rr = rr;
    sc->rho = rho;
    sc->rho_prev = rho_prev;
    sc->pq = pq;
    sc->alpha = alpha;
    sc->beta = beta;
    sc->iter = iter;
    sc->done = 1;
    sc->converged = converged;
    sc->breakdown = breakdown;
  }
}

}  // namespace b200spai
