// Whole PCG solve of a SMALL system (<= 131 072 rows) in one co-resident kernel -- second generation of pcg_small.cuh
// (same three phases per iteration, same arithmetic; BASELINE configs C1-C4 and the C2 batches run here):
//
//   P1   [x += alpha p ; r' = r - alpha q ; ||r'||^2]  and  t = G^T r'   (r' of a neighbour row recomputed on the fly)
//   --- all-reduce: ||r'||^2 -> iteration count, history, convergence / max_iter test (scipy cg order) ---
//   P2   z = G t + eps r' [/ d] ; rho = r'.z
//   --- all-reduce: rho -> beta ---
//   P3   p' = z + beta p  and  q = A p'  (p' of a neighbour column recomputed on the fly) ; p'.q
//   --- all-reduce: p.q -> alpha (breakdown if <= 0) ---
//
// What changed against the thread-per-row kernel (measured there: 20 us per iteration at 6 k rows, 33-72 us on the
// irregular synthetic matrices with rows of up to 118 entries -- three latency chains of sequential row walks):
//
//   * a row is walked by L = 4 / 8 / 16 / 32 lanes (chosen from the mean row length): all entries of a pass and their
//     gathered vector entries are in flight at once (two L2 round trips per row, whatever its length), the products are
//     rounded per lane, and the row sum is then formed IN ENTRY ORDER by passing the products through a shuffle
//     (sum = sum + p_k, one rounded add per entry) -- the bits of scipy's sequential csr_matvec, like every other SpMV of
//     the library;
//   * no cooperative-groups grid.sync + separate partial-sum read: a CTA publishes its two partial sums as generation-
//     tagged LL words (value and flag in one 16-byte store) and EVERY CTA polls all slots itself -- one L2 round trip from
//     the last arrival to the totals, summed in a fixed order so that all CTAs hold identical scalars and take identical
//     branches without a broadcast;
//   * the CTA's rows of G^T, G and A (rowptr, columns, fp64 values) are copied into shared memory once and read from there
//     in every iteration: the only L2 round trip left in a phase is the gather of the vector entries (the row owner's
//     own operands are requested before the row walk and overlap it);
//   * waits give up after spin_limit clock ticks (breakdown = 3) instead of hanging the GPU.
#pragma once
#include "pcg_kernels.cuh"

namespace b200spai {

struct Small2Args {
  CsrView A, G, GT;
  const double* b;
  double* x;
  double *r0, *r1, *p0, *p1, *z, *q, *t;
  const double* d;  // diagonal of A (diagonal / scaled SPAI) or nullptr
  double eps;
  int precond;
  PcgScalars* sc;
  unsigned long long* slots;  // [2][gridDim.x][4] LL words, zeroed before the launch
  double* hist;
  long long spin_limit;
  int smem_bytes;  // dynamic shared memory of the launch
  // Batched launch (b200spai_pcg_solve_batch): A, G, G^T are block diagonal, block s = rows [row0, row0 + nrows) = one
  // independent system, solved by the CTA group [cta0, cta0 + ncta) with its own scalars sc[s], all-reduce slots and
  // (optionally) history hist + s * hist_stride.  nullptr: one system, the whole grid.
  const struct S2Sys* sys;
  int n_sys;
  long long hist_stride;
};

struct S2Sys {
  int row0, nrows, cta0, ncta;
};

constexpr int S2_THREADS = 256;
constexpr int S2_MAX_CTAS = 512;  // 2 * 512 * 4 words = 32 KB of the 64 KB partials buffer

__device__ __forceinline__ void s2_ll_store(unsigned long long* slot, double v, unsigned long long tag32) {
  const unsigned long long bits = (unsigned long long)__double_as_longlong(v);
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(slot), "l"(((bits & 0xffffffffull) << 32) | tag32),
               "l"((bits & 0xffffffff00000000ull) | tag32) : "memory");
}
__device__ __forceinline__ bool s2_ll_load(const unsigned long long* slot, unsigned long long tag32, double& v) {
  unsigned long long w0, w1;
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(slot) : "memory");
  v = __longlong_as_double((long long)((w1 & 0xffffffff00000000ull) | (w0 >> 32)));
  return (w0 & 0xffffffffull) == tag32 && (w1 & 0xffffffffull) == tag32;
}

// (v0, v1) summed over the grid; identical bits in every thread of every CTA.  Doubles as the grid barrier: everything
// the CTAs wrote before the call is visible (through L2: __ldcg / ld.relaxed) after it.  false: a CTA never arrived.
struct S2Group {  // the CTA group that solves one system
  unsigned long long* slots;  // [2][G][4]
  unsigned G, bid;
  long long spin_limit;
};

__device__ __forceinline__ bool s2_allreduce(const S2Group& a, unsigned long long& gen, double v0, double v1, double* red,
                                             double& t0, double& t1) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned G = a.G;
  gen += 1;
  unsigned long long* slot = a.slots + (size_t)(gen & 1ull) * G * 4;
  const unsigned long long g32 = gen & 0xffffffffull;
  // ---- this CTA's partials: warp tree, then the 8 warps in index order ----
  v0 = warp_sum(v0);
  v1 = warp_sum(v1);
  __syncthreads();  // protects `red`; every thread's vector writes precede thread 0's fence
  if (lane == 0) {
    red[warp] = v0;
    red[8 + warp] = v1;
  }
  __syncthreads();
  if (tid < 2) {
    const double* w = red + 8 * tid;
    const double c = ((w[0] + w[1]) + (w[2] + w[3])) + ((w[4] + w[5]) + (w[6] + w[7]));
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
    s2_ll_store(slot + 4 * a.bid + 2 * tid, c, g32);
  }
  // ---- every CTA fetches every CTA's partials (the poll is the barrier) ----
  double s0 = 0.0, s1 = 0.0;
  bool okp = true;
  const long long c0k = clock64();
  for (unsigned i = tid; i < G; i += S2_THREADS) {
    double b0 = 0.0, b1 = 0.0;
    while (!(s2_ll_load(slot + 4 * i, g32, b0) & s2_ll_load(slot + 4 * i + 2, g32, b1)))
      if (clock64() - c0k > a.spin_limit) {
        okp = false;
        break;
      }
    s0 += b0;
    s1 += b1;
  }
  asm volatile("fence.acq_rel.gpu;" ::: "memory");
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  const int timed_out = __syncthreads_or(!okp);  // also protects `red`
  if (lane == 0) {
    red[16 + warp] = s0;
    red[24 + warp] = s1;
  }
  __syncthreads();
  const double* w0 = red + 16;
  const double* w1 = red + 24;
  t0 = ((w0[0] + w0[1]) + (w0[2] + w0[3])) + ((w0[4] + w0[5]) + (w0[6] + w0[7]));
  t1 = ((w1[0] + w1[1]) + (w1[2] + w1[3])) + ((w1[4] + w1[5]) + (w1[6] + w1[7]));
  return !timed_out;
}

// A CTA's rows of one matrix: rowptr / col / val indexed with GLOBAL row and entry numbers, the pointers biased so that they
// address either the CTA's copy in shared memory or the arrays in global memory (generic addresses, one code path).
struct S2View {
  const int* rowptr;
  const int* col;
  const double* val;
};

// Copies rows [r_lo, r_hi) of M into shared memory at `sp` when they fit into `room` bytes (all threads of the CTA);
// returns the view and advances sp / room.  The staged copy is what every iteration of the solve reads: no L2 round trip
// for the matrix, only for the gathered vector entries.
__device__ __forceinline__ S2View s2_stage(const CsrView& M, int r_lo, int r_hi, unsigned char*& sp, int& room) {
  S2View g{M.rowptr, M.col, M.val};
  if (r_hi <= r_lo || M.rowptr == nullptr) return g;
  const int s = __ldg(M.rowptr + r_lo), e = __ldg(M.rowptr + r_hi);
  const int rows = r_hi - r_lo, cnt = e - s;
  const int bytes = ((cnt * 8 + 15) & ~15) + ((cnt * 4 + 15) & ~15) + (((rows + 1) * 4 + 15) & ~15);
  if (bytes > room) return g;  // this CTA's share does not fit: it keeps reading global memory (slower, same bits)
  double* sv = reinterpret_cast<double*>(sp);
  int* sc = reinterpret_cast<int*>(sp + ((cnt * 8 + 15) & ~15));
  int* sr = sc + (((cnt * 4 + 15) & ~15) >> 2);
  for (int k = threadIdx.x; k < cnt; k += blockDim.x) {
    sv[k] = __ldg(M.val + s + k);
    sc[k] = __ldg(M.col + s + k);
  }
  for (int k = threadIdx.x; k <= rows; k += blockDim.x) sr[k] = __ldg(M.rowptr + r_lo + k);
  sp += bytes;
  room -= bytes;
  return S2View{sr - r_lo, sc - s, sv - s};
}

// One row of M times a vector whose entry c is produced by `entry(c)`, walked by the L = 1 << LG lanes of a group (lane j
// of the group takes entries j, j + L, ...); called by all 32 lanes of a warp together (`valid` = this group has a row).
// Eight passes of loads (16 with one lane per row) are in flight at once; the sum is formed in entry order (rounded
// product, rounded add).  Every lane of the group returns the row sum.
// (s, e) of a group's row, requested one row ahead of its use (the row loop prefetches the next pair while it walks the
// current row: with the matrices in global memory that takes one of three dependent L2 round trips off every pass)
struct S2Range {
  int s, e;
};
__device__ __forceinline__ S2Range s2_range(const S2View& M, int row, bool valid) {
  S2Range r{0, 0};
  if (valid) {
    r.s = M.rowptr[row];
    r.e = M.rowptr[row + 1];
  }
  return r;
}

template <int LG, class Entry>
__device__ __forceinline__ double s2_row(const S2View& M, const S2Range rg, int j, Entry entry) {
  constexpr int L = 1 << LG;
  const int s = rg.s, e = rg.e;
  const int len = e - s;
  const int npass = __reduce_max_sync(0xffffffffu, (len + L - 1) >> LG);
  double sum = 0.0;
  constexpr int U = (LG == 0) ? 16 : 8;  // thread per row: a 15-entry mesh row in one round
  for (int ps = 0; ps < npass; ps += U) {
    int c[U];
    double v[U], pr[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int k = s + ((ps + u) << LG) + j;
      const bool in = k < e;
      c[u] = in ? M.col[k] : -1;
      v[u] = in ? M.val[k] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) pr[u] = (c[u] >= 0) ? __dmul_rn(v[u], entry(c[u])) : 0.0;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (ps + u < npass) {  // warp-uniform
        const int m = len - ((ps + u) << LG);  // entries of this pass in my row (<= 0: none)
#pragma unroll
        for (int i = 0; i < L; ++i) {
          const double o = (L > 1) ? __shfl_sync(0xffffffffu, pr[u], i, L) : pr[u];
          if (i < m) sum = __dadd_rn(sum, o);
        }
      }
    }
  }
  return sum;
}

template <int LG>
__global__ void __launch_bounds__(S2_THREADS, 2) pcg_small2_kernel(const Small2Args a) {
  constexpr int L = 1 << LG;
  constexpr int NG = S2_THREADS >> LG;  // row groups per CTA
  __shared__ double red[32];
  extern __shared__ __align__(16) unsigned char s2_smem[];
  const int tid = threadIdx.x;
  // which system, which CTA of its group
  int row0 = 0, n = a.A.n_rows, cta0 = 0, G = gridDim.x, sys_id = 0;
  if (a.sys != nullptr) {
    __shared__ int s_sys;
    if (tid == 0) {
      int lo = 0, hi = a.n_sys - 1;  // groups are consecutive CTA ranges
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (a.sys[mid].cta0 <= (int)blockIdx.x) lo = mid;
        else hi = mid - 1;
      }
      s_sys = lo;
    }
    __syncthreads();
    sys_id = s_sys;
    const S2Sys d = a.sys[sys_id];
    row0 = d.row0;
    n = d.nrows;
    cta0 = d.cta0;
    G = d.ncta;
  }
  const int bid = (int)blockIdx.x - cta0;
  const int chunk = (n + G - 1) / G;
  const int r_lo = row0 + min(n, bid * chunk), r_hi = row0 + min(n, bid * chunk + chunk);
  const int grp = tid >> LG, j = tid & (L - 1);
  const bool spai = a.precond >= B200SPAI_PRECOND_EXT_SPAI;
  const S2Group grpc{a.slots + (size_t)cta0 * 8, (unsigned)G, (unsigned)bid, a.spin_limit};
  double* const hist = a.hist ? a.hist + (long long)sys_id * a.hist_stride : nullptr;
  // the CTA's rows of G^T, G and A live in shared memory for the whole solve (when they fit)
  unsigned char* sp = s2_smem;
  int room = a.smem_bytes;
  const S2View vGT = spai ? s2_stage(a.GT, r_lo, r_hi, sp, room) : S2View{nullptr, nullptr, nullptr};
  const S2View vG = spai ? s2_stage(a.G, r_lo, r_hi, sp, room) : S2View{nullptr, nullptr, nullptr};
  const S2View vA = s2_stage(a.A, r_lo, r_hi, sp, room);
  const bool scaled = a.precond == B200SPAI_PRECOND_EXT_SPAI_SCALED;
  const bool lead = (bid == 0 && tid == 0);
  PcgScalars* sc = a.sc + sys_id;
  const double rtol = sc->rtol;
  const int max_iter = sc->max_iter;
  unsigned long long gen = 0;
  double t0 = 0.0, t1 = 0.0;

  // ---- init: r = b, x = 0, ||b|| ---------------------------------------------------------------------
  double acc = 0.0;
  for (int i = r_lo + tid; i < r_hi; i += S2_THREADS) {
    const double v = a.b[i];
    a.r0[i] = v;
    a.x[i] = 0.0;
    acc += v * v;
  }
  bool ok = s2_allreduce(grpc, gen, acc, 0.0, red, t0, t1);
  const double bb = t0;
  const double bnorm = sqrt(bb);
  const double atol = fmax(0.0, rtol * bnorm);
  double rr = bb, rho = 0.0, rho_prev = 0.0, pq = 0.0, alpha = 0.0, beta = 0.0;
  int iter = 0, done = 0, converged = 0, breakdown = 0;
  if (!ok) {
    done = 1;
    breakdown = 3;
  } else {
    if (lead && hist) hist[0] = bnorm;
    if (bnorm == 0.0) {
      done = 1;
      converged = 1;
    } else if (!(bnorm == bnorm)) {
      done = 1;
      breakdown = 1;
    }
  }
  double *r = a.r0, *rn = a.r1, *p = a.p0, *pn = a.p1;
  bool pending = false;  // an (alpha, p, q) update of x and r is waiting to be applied in P1

  while (!done) {
    // ---- P1 -------------------------------------------------------------------------------------------
    double rr_part = 0.0, rz_part = 0.0;
    if (spai) {
      S2Range nxt = s2_range(vGT, r_lo + grp, r_lo + grp < r_hi);
      for (int base = r_lo; base < r_hi; base += NG) {
        const int i = base + grp;
        const bool valid = i < r_hi;
        const bool mine = valid && j == 0;
#ifdef B200SPAI_S2_NO_PREFETCH
        const S2Range cur = s2_range(vGT, i, valid);
#else
        const S2Range cur = nxt;
        nxt = s2_range(vGT, i + NG, i + NG < r_hi);
#endif
        // the row owner's operands are requested before the row walk: their L2 round trip overlaps the gathers
        double xi = 0.0, pi = 0.0, ri = 0.0, qi = 0.0, di = 1.0;
        if (mine) {
          if (pending) {
            xi = __ldcg(a.x + i);
            pi = __ldcg(p + i);
            ri = __ldcg(r + i);
            qi = __ldcg(a.q + i);
          }
          if (scaled) di = __ldg(a.d + i);
        }
        double tv;
        if (pending) {
          const double* rq = a.q;
          const double al = alpha;
          tv = s2_row<LG>(vGT, cur, j, [r, rq, al](int c) { return __dsub_rn(__ldcg(r + c), __dmul_rn(al, __ldcg(rq + c))); });
        } else {
          tv = s2_row<LG>(vGT, cur, j, [r](int c) { return __ldcg(r + c); });
        }
        if (mine) {
          if (pending) {
            a.x[i] = __dadd_rn(xi, __dmul_rn(alpha, pi));
            const double rv = __dsub_rn(ri, __dmul_rn(alpha, qi));
            rn[i] = rv;
            rr_part += rv * rv;
          }
          a.t[i] = scaled ? tv / di : tv;
        }
      }
    } else {
      for (int i = r_lo + tid; i < r_hi; i += S2_THREADS) {
        double rv = r[i];
        if (pending) {
          a.x[i] = __dadd_rn(a.x[i], __dmul_rn(alpha, p[i]));
          rv = __dsub_rn(rv, __dmul_rn(alpha, a.q[i]));
          rn[i] = rv;
          rr_part += rv * rv;
        }
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
    ok = s2_allreduce(grpc, gen, rr_part, rz_part, red, t0, t1);  // first SPAI iteration: only t has to be complete
    if (!ok) {
      breakdown = 3;
      break;
    }
    if (pending) {  // the update is complete: count it, test it (scipy: callback, then the test at the loop top)
      if (!(t0 == t0)) {
        breakdown = 1;
        break;
      }
      rr = t0;
      rho_prev = rho;
      iter += 1;
      const double nrm = sqrt(t0);
      if (lead && hist) hist[iter] = nrm;
      if (nrm < atol) {
        converged = 1;
        break;
      }
      if (iter >= max_iter) break;
    }
    if (!spai) rho = t1;
    // ---- P2 (SPAI): z = G t + eps r [/ d], rho = r.z ------------------------------------------------------
    if (spai) {
      double part = 0.0;
      const double* tt = a.t;
      S2Range nxt = s2_range(vG, r_lo + grp, r_lo + grp < r_hi);
      for (int base = r_lo; base < r_hi; base += NG) {
        const int i = base + grp;
        const bool valid = i < r_hi;
        const bool mine = valid && j == 0;
#ifdef B200SPAI_S2_NO_PREFETCH
        const S2Range cur = s2_range(vG, i, valid);
#else
        const S2Range cur = nxt;
        nxt = s2_range(vG, i + NG, i + NG < r_hi);
#endif
        double rv = 0.0, di = 1.0;
        if (mine) {
          rv = __ldcg(r + i);
          if (scaled) di = __ldg(a.d + i);
        }
        const double gv = s2_row<LG>(vG, cur, j, [tt](int c) { return __ldcg(tt + c); });
        if (mine) {
          double e = a.eps * rv;
          if (scaled) e = e / di;
          const double zv = gv + e;
          a.z[i] = zv;
          part += rv * zv;
        }
      }
      ok = s2_allreduce(grpc, gen, part, 0.0, red, t0, t1);
      if (!ok) {
        breakdown = 3;
        break;
      }
      rho = t0;
    }
    const bool first = (iter == 0);
    beta = first ? 0.0 : rho / rho_prev;
    // ---- P3: p' = z + beta p, q = A p', p'.q --------------------------------------------------------------
    double part = 0.0;
    {
      const double* zz = a.z;
      const double be = beta;
      S2Range nxt = s2_range(vA, r_lo + grp, r_lo + grp < r_hi);
      for (int base = r_lo; base < r_hi; base += NG) {
        const int i = base + grp;
        const bool valid = i < r_hi;
        const bool mine = valid && j == 0;
#ifdef B200SPAI_S2_NO_PREFETCH
        const S2Range cur = s2_range(vA, i, valid);
#else
        const S2Range cur = nxt;
        nxt = s2_range(vA, i + NG, i + NG < r_hi);
#endif
        double zi = 0.0, pi = 0.0;
        if (mine) {
          zi = __ldcg(zz + i);
          if (!first) pi = __ldcg(p + i);
        }
        double qv;
        if (first) qv = s2_row<LG>(vA, cur, j, [zz](int c) { return __ldcg(zz + c); });
        else qv = s2_row<LG>(vA, cur, j, [zz, p, be](int c) { return __dadd_rn(__dmul_rn(__ldcg(p + c), be), __ldcg(zz + c)); });
        if (mine) {
          const double pv = first ? zi : __dadd_rn(__dmul_rn(pi, beta), zi);
          pn[i] = pv;
          a.q[i] = qv;
          part += pv * qv;
        }
      }
    }
    ok = s2_allreduce(grpc, gen, part, 0.0, red, t0, t1);
    if (!ok) {
      breakdown = 3;
      break;
    }
    pq = t0;
    {
      double* tmp = p;
      p = pn;
      pn = tmp;
    }
    if (!(pq > 0.0) || !(rho == rho)) {  // p.q <= 0 or NaN: A or M not SPD
      breakdown = 1;
      alpha = 0.0;
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
