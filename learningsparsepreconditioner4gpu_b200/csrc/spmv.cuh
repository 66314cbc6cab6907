// CSR SpMV kernels (fp64 values, int32 indices) with fused PCG epilogues.
//
// Two kernels, both persistent (grid = SM count x resident CTAs) and both deterministic:
//
//  * spmv_vector_kernel  -- a sub-warp of VEC lanes per row, direct global loads, two row batches in
//    flight per warp.  Robust baseline; used for tiny matrices and very irregular ones.
//  * spmv_stream_kernel  -- the B200 path.  Rows are pre-partitioned (once per matrix) into row
//    blocks of <= CAP non-zeros.  CSR values / column indices / row pointers of a block are
//    *contiguous* in HBM, so each block is staged into shared memory with three 1-D TMA bulk copies
//    (cp.async.bulk + mbarrier complete_tx) by a producer warp, STAGES blocks deep, with an L2
//    evict-first policy so the streamed matrix does not displace the vectors that are gathered out
//    of L2.  Consumer warps process a staged block row-parallel (thread = row) straight out of
//    shared memory and hand the slot back through an "empty" mbarrier.  Blocks are dealt to CTAs
//    round-robin, so all CTAs sweep the matrix as one wavefront and the gathered window of x
//    stays L2/L1 resident.  Values are streamed as fp32 when that is lossless (see below).
//
// Epilogues (MODE):   PLAIN      y = Ax
//                     SCALE      y = (Ax) / d                          t = (G^T r)/d   (scaled SPAI)
//                     AXPY_DOT   y = Ax + eps*aux [/ d];  dot = aux.y    z = G t + eps r; rho = r.z
//                     DOT        y = Ax;  dot = x.y                     q = A p; pq = p.q
// The dot is reduced deterministically (fixed per-CTA order, then the last CTA sums the per-CTA
// partials in index order) and the last CTA also performs the scalar update (beta / alpha).
#pragma once
#include "common.cuh"

namespace b200spai {

struct PcgScalars {
  double rho, rho_prev, pq, rr, alpha, beta, bnorm, atol, rtol;
  int iter, done, converged, breakdown, max_iter;
  int use_cond;
  unsigned long long cond_handle;
  // row-partitioned solve (dist.cu): reductions stop at the local total, an all-reduce kernel finishes them
  int dist;
  int have_rr;        // dist: ||r||^2 of the update just made is waiting in local_rr (all-reduced together with rho)
  double local_total;
  double local_rr;
};

struct CsrView {
  int n_rows;
  const int* rowptr;
  const int* col;
  const double* val;
};

enum { SPMV_PLAIN = 0, SPMV_SCALE = 1, SPMV_AXPY_DOT = 2, SPMV_DOT = 3 };
enum { FINAL_NONE = 0, FINAL_RHO = 1, FINAL_PQ = 2 };

struct SpmvArgs {
  const double* x;
  double* y;
  const double* aux;  // AXPY_DOT: the vector added with eps and dotted with y;  DOT: x restricted to this matrix's rows
  const double* d;    // optional diagonal (divide), SCALE and AXPY_DOT
  double eps;
  double* partials;
  unsigned* ticket;
  PcgScalars* sc;
  int final_op;
};

__device__ __forceinline__ void pcg_final_scalar(PcgScalars* sc, int final_op, double total) {
  if (sc->dist) {  // row-partitioned solve: the all-reduce kernel that follows applies the update to the global total
    sc->local_total = total;
    return;
  }
  if (final_op == FINAL_RHO) {
    sc->rho = total;
    sc->beta = (sc->iter > 0) ? total / sc->rho_prev : 0.0;
  } else if (final_op == FINAL_PQ) {
    sc->pq = total;
    if (!(total > 0.0) || !(sc->rho == sc->rho)) {  // p.q <= 0 or NaN: not SPD / breakdown
      sc->breakdown = 1;
      sc->alpha = 0.0;
    } else {
      sc->alpha = sc->rho / total;
    }
  }
}

template <int MODE>
__device__ __forceinline__ double spmv_prefetch_aux(const SpmvArgs& a, int row) {
  if (MODE == SPMV_AXPY_DOT || MODE == SPMV_DOT) return a.aux[row];
  return 0.0;
}

template <int MODE>
__device__ __forceinline__ void spmv_row_epilogue(const SpmvArgs& a, int row, double sum, double auxv, double& dot) {
  double yv = sum;
  if (MODE == SPMV_SCALE) {
    yv = sum / a.d[row];
  } else if (MODE == SPMV_AXPY_DOT) {
    double e = a.eps * auxv;
    if (a.d) e = e / a.d[row];
    yv = sum + e;
    dot += auxv * yv;
  } else if (MODE == SPMV_DOT) {
    dot += auxv * yv;
  }
  a.y[row] = yv;
}

template <int MODE>
__device__ __forceinline__ void spmv_finish_dot(const SpmvArgs& a, double dot, double* red) {
  if (MODE == SPMV_AXPY_DOT || MODE == SPMV_DOT) {
    double cta = block_sum(dot, red);
    double total;
    if (grid_sum_last(cta, a.partials, a.ticket, red, total)) {
      if (threadIdx.x == 0 && a.sc) pcg_final_scalar(a.sc, a.final_op, total);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Kernel 1: sub-warp per row
// ------------------------------------------------------------------------------------------------
template <int VEC, int MODE>
__global__ void __launch_bounds__(256) spmv_vector_kernel(CsrView A, SpmvArgs a) {
  __shared__ double red[34];
  if (a.sc && a.sc->done) return;
  constexpr int RPW = 32 / VEC;  // rows per warp per batch
  const int lane = threadIdx.x & 31;
  const int sub = lane / VEC, sl = lane % VEC;
  const long long warp_g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long stride = nwarps * RPW;
  double dot = 0.0;
  for (long long rb = warp_g * RPW; rb < A.n_rows; rb += 2 * stride) {
    const long long row0 = rb + sub, row1 = rb + stride + sub;
    const bool v0 = row0 < A.n_rows, v1 = row1 < A.n_rows;
    int s0 = 0, e0 = 0, s1 = 0, e1 = 0;
    if (v0) { s0 = __ldg(A.rowptr + row0); e0 = __ldg(A.rowptr + row0 + 1); }
    if (v1) { s1 = __ldg(A.rowptr + row1); e1 = __ldg(A.rowptr + row1 + 1); }
    double aux0 = 0.0, aux1 = 0.0;
    if (sl == 0) {
      if (v0) aux0 = spmv_prefetch_aux<MODE>(a, (int)row0);
      if (v1) aux1 = spmv_prefetch_aux<MODE>(a, (int)row1);
    }
    double sum0 = 0.0, sum1 = 0.0;
    int k0 = s0 + sl, k1 = s1 + sl;
    while (k0 < e0 || k1 < e1) {
      double p0 = 0.0, p1 = 0.0;
      if (k0 < e0) p0 = __ldg(A.val + k0) * __ldg(a.x + __ldg(A.col + k0));
      if (k1 < e1) p1 = __ldg(A.val + k1) * __ldg(a.x + __ldg(A.col + k1));
      sum0 += p0;
      sum1 += p1;
      k0 += VEC;
      k1 += VEC;
    }
#pragma unroll
    for (int o = VEC / 2; o > 0; o >>= 1) {
      sum0 += __shfl_xor_sync(0xffffffffu, sum0, o);
      sum1 += __shfl_xor_sync(0xffffffffu, sum1, o);
    }
    if (sl == 0) {
      if (v0) spmv_row_epilogue<MODE>(a, (int)row0, sum0, aux0, dot);
      if (v1) spmv_row_epilogue<MODE>(a, (int)row1, sum1, aux1, dot);
    }
  }
  spmv_finish_dot<MODE>(a, dot, red);
}

// ------------------------------------------------------------------------------------------------
// Kernel 2: TMA-staged, warp-specialised, row-parallel streaming kernel
// ------------------------------------------------------------------------------------------------
// One producer warp (one elected lane) keeps STAGES row blocks in flight with 1-D bulk copies
// (values, column indices, row pointers -> shared memory, mbarrier complete_tx, L2 evict-first);
// NCW consumer warps take a staged block ROW-PARALLEL: thread = row, walking its entries out of
// shared memory.  Adjacent rows of a mesh matrix have adjacent column indices, so the gather
// x[col] of one k-step of a warp touches 2-3 128-byte lines instead of the ~16 an nnz-parallel
// sweep touches (the L1 pipeline takes about one cycle per line, which bounded the previous
// version at ~1/3 of HBM peak).  Slots are recycled through per-slot "empty" mbarriers, so no
// CTA-wide barrier sits in the loop and the consumer warps drift freely over the stages.
//
// The matrix stream is compressed losslessly, 12 -> 6 bytes per entry when both apply:
//  * VT = float when every value of the matrix is exactly representable in fp32 (true for A and G
//    of this path: both are widened fp32 tensors) -- values are widened on load, arithmetic stays
//    fp64, results are bit-identical;
//  * CT = uint16_t when every within-row column step fits 16 bits (any matrix with < 65536 columns,
//    and banded mesh matrices of any size): entry k stores col[k] - col[k-1] (0 for the first entry
//    of a row, whose column travels in a 4-byte-per-row side array).  The thread that owns the row
//    walks it sequentially anyway, so decoding is one integer add per entry.
#ifdef B200SPAI_PROBE
static __device__ unsigned long long g_probe[8];
#define PROBE_T(v) const long long v = clock64()
#define PROBE_ADD(i, val) atomicAdd(&g_probe[i], (unsigned long long)(val))
#else
#define PROBE_T(v)
#define PROBE_ADD(i, val)
#endif

#ifndef B200SPAI_PART_W
#define B200SPAI_PART_W 8
#endif
constexpr int PART_W = B200SPAI_PART_W;  // row-block partition key(r) = rowptr[r] + PART_W * r: a block holds <= TARGET / PART_W rows

template <int CAP, typename VT, typename CT>
struct StreamCfg {
  static constexpr bool DELTA = sizeof(CT) == 2;
  static constexpr int SLACK = CAP / 8;               // rows up to SLACK long never overflow a block
  static constexpr int TARGET = CAP - SLACK;          // partition target on key(r)
  static constexpr int RCAP = TARGET / PART_W + 2;    // max rows per block + 1
  static constexpr int VALS = CAP + 8;                // staged values (alignment slack both ends)
  static constexpr int COLS = CAP + 32;               // + the 16-entry chunk over-read
  static constexpr int RPTR = ((RCAP + 8 + 3) / 4) * 4;
  static constexpr int VALS_BYTES = VALS * (int)sizeof(VT);
  static constexpr int COLS_BYTES = ((COLS * (int)sizeof(CT) + 15) / 16) * 16;
  static constexpr int STAGE_BYTES = VALS_BYTES + COLS_BYTES + RPTR * 4 * (DELTA ? 2 : 1);
  static constexpr int smem_bytes(int stages) { return stages * STAGE_BYTES + stages * 16 /*hdr*/ + stages * 16 /*bars*/ + 34 * 8; }
};

template <int CAP, int STAGES, int NCW, typename VT, typename CT, int MODE>
__global__ void __launch_bounds__((NCW + 1) * 32)
    spmv_stream_kernel(CsrView A, const VT* __restrict__ Aval, const CT* __restrict__ Acol, const int* __restrict__ Afirst,
                       const int4* __restrict__ blk, int nblk, SpmvArgs a) {
  using C = StreamCfg<CAP, VT, CT>;
  constexpr bool DELTA = C::DELTA;
  constexpr int NCT = NCW * 32;  // consumer threads
  constexpr int CA = 16 / (int)sizeof(CT);  // column entries per 16 bytes (bulk-copy alignment unit)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  if (a.sc && a.sc->done) return;
  unsigned char* sp = smem_raw;
  VT* s_vals = reinterpret_cast<VT*>(sp);                 sp += STAGES * C::VALS_BYTES;          // [STAGES][VALS]
  CT* s_cols = reinterpret_cast<CT*>(sp);                 sp += STAGES * C::COLS_BYTES;          // [STAGES][COLS]
  int* s_rptr = reinterpret_cast<int*>(sp);               sp += STAGES * C::RPTR * 4;            // [STAGES][RPTR]
  int* s_first = reinterpret_cast<int*>(sp);              sp += DELTA ? STAGES * C::RPTR * 4 : 0;  // [STAGES][RPTR]
  int4* s_hdr = reinterpret_cast<int4*>(sp);              sp += STAGES * 16;                     // [STAGES]
  uint64_t* s_full = reinterpret_cast<uint64_t*>(sp);     sp += STAGES * 8;                      // [STAGES]
  uint64_t* s_empty = reinterpret_cast<uint64_t*>(sp);    sp += STAGES * 8;                      // [STAGES]
  double* red = reinterpret_cast<double*>(sp);                                                   // [34]

  const int tid = threadIdx.x;
  const int G = gridDim.x;
  const int first = blockIdx.x;
  const int n_items = (first < nblk) ? (nblk - first + G - 1) / G : 0;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(s_full + s, 1);
      mbar_init(s_empty + s, NCW);
    }
    mbar_fence_init();
  }
  __syncthreads();

  double dot = 0.0;
  if (tid >= NCT) {
    // ---------------- producer warp ----------------
    if (tid == NCT) {
      const uint64_t policy = policy_evict_first();
      int4 info = make_int4(0, 0, 0, 0);
      if (n_items > 0) info = __ldg(blk + first);
      for (int j = 0; j < n_items; ++j) {
        const int slot = j % STAGES;
        PROBE_T(pt0);
        if (j >= STAGES) mbar_wait(s_empty + slot, (uint32_t)(((j / STAGES) - 1) & 1));
        PROBE_T(pt1);
        PROBE_ADD(2, pt1 - pt0);
        const int4 cur = info;
        if (j + 1 < n_items) info = __ldg(blk + first + (long long)(j + 1) * G);
        s_hdr[slot] = cur;
        const int r0 = cur.x, lo = cur.y, r1 = cur.z, hi = cur.w;
        const int cnt = hi - lo;
        if (cnt > CAP || cnt <= 0) {  // over-long row or all-empty rows: consumers read global memory
          mbar_arrive(s_full + slot);
          continue;
        }
        const int v_lo = lo & ~3;                       // values: 4-byte entries, 16-byte aligned start
        const int vn = (hi - v_lo + 3) & ~3;
        const int c_lo = lo & ~(CA - 1);                // columns: CA entries per 16 bytes
        const int cn = (hi - c_lo + CA - 1) & ~(CA - 1);
        const int ra = r0 & ~3;
        const int nr4 = (r1 + 1 - ra + 3) & ~3;
        // VT = double: 8-byte entries, v_lo & ~3 is still 16-byte aligned
        mbar_arrive_expect_tx(s_full + slot, (uint32_t)(vn * (int)sizeof(VT) + cn * (int)sizeof(CT) + nr4 * 4 * (DELTA ? 2 : 1)));
        bulk_g2s(s_vals + slot * C::VALS, Aval + v_lo, (uint32_t)(vn * sizeof(VT)), s_full + slot, policy);
        bulk_g2s(reinterpret_cast<unsigned char*>(s_cols) + slot * C::COLS_BYTES, Acol + c_lo, (uint32_t)(cn * sizeof(CT)),
                 s_full + slot, policy);
        bulk_g2s(s_rptr + slot * C::RPTR, A.rowptr + ra, (uint32_t)nr4 * 4, s_full + slot, policy);
        if (DELTA) bulk_g2s(s_first + slot * C::RPTR, Afirst + ra, (uint32_t)nr4 * 4, s_full + slot, policy);
      }
    }
  } else {
    // ---------------- consumer warps ----------------
    const int lane = tid & 31, warp = tid >> 5;
    const double* __restrict__ x = a.x;
    for (int i = 0; i < n_items; ++i) {
      const int slot = i % STAGES;
      PROBE_T(ct0);
      mbar_wait(s_full + slot, (uint32_t)((i / STAGES) & 1));
      PROBE_T(ct1);
      if (tid == 0) PROBE_ADD(0, ct1 - ct0);
      const int4 info = s_hdr[slot];
      const int r0 = info.x, lo = info.y, r1 = info.z, hi = info.w;
      const int cnt = hi - lo;
      if (cnt > 0 && cnt <= CAP) {
        const VT* vals = s_vals + slot * C::VALS + (lo & 3);
        const CT* cols = reinterpret_cast<const CT*>(reinterpret_cast<const unsigned char*>(s_cols) + slot * C::COLS_BYTES) +
                         (lo & (CA - 1));
        const int* rptr = s_rptr + slot * C::RPTR + (r0 & 3);
        const int* rfirst = s_first + slot * C::RPTR + (r0 & 3);
        for (int rr = tid; rr < r1 - r0; rr += NCT) {
          const int row = r0 + rr;
          const double auxv = spmv_prefetch_aux<MODE>(a, row);
          int k = rptr[rr] - lo;
          const int e = rptr[rr + 1] - lo;
          int cprev = DELTA ? rfirst[rr] : 0;
          // Row sum in scipy's csr_matvec order and arithmetic (sequential, product rounded, then added -- no
          // fma contraction), so y is bit-identical to `A @ x` of the reference's CPU path.  The gathers of a
          // 16-entry chunk go out together as asynchronous copies (one exposed memory latency per chunk, and a
          // mesh row is one chunk); only the adds form a dependent chain.
          double sum = 0.0;
#pragma unroll 1
          for (; k < e; k += 16) {
            int c[16];
            double pr[16];
            // entries past the row end re-read the last valid column (unconditional loads) and multiply by +0
#pragma unroll
            for (int u = 0; u < 16; ++u) {
              const int raw = (int)cols[k + u];
              if (DELTA) {
                cprev += (k + u < e) ? raw : 0;
              } else {
                cprev = (k + u < e) ? raw : cprev;
              }
              c[u] = cprev;
            }
            // the 16 products are independent of each other, so their gathers go out back to back ...
#pragma unroll
            for (int u = 0; u < 16; ++u) pr[u] = __dmul_rn((k + u < e) ? (double)vals[k + u] : 0.0, __ldg(x + c[u]));
            // ... and only the adds form a dependent chain (sum + (+-0) == sum bit for bit: sum starts at +0 and
            // is therefore never -0)
#pragma unroll
            for (int u = 0; u < 16; ++u) sum = __dadd_rn(sum, pr[u]);
          }
          spmv_row_epilogue<MODE>(a, row, sum, auxv, dot);
        }
      } else {
        // a row longer than the stage capacity (or a block of empty rows): warp per row from global memory.  Lanes
        // fetch 32 products at a time, then the warp adds them in entry order, so these rows too carry scipy's bits.
        for (int row = r0 + warp; row < r1; row += NCW) {
          const int s = __ldg(A.rowptr + row), e = __ldg(A.rowptr + row + 1);
          double sum = 0.0;
          for (int k0 = s; k0 < e; k0 += 32) {
            const int k = k0 + lane;
            const double p = (k < e) ? __dmul_rn((double)__ldg(Aval + k), __ldg(x + __ldg(A.col + k))) : 0.0;
            const int m = min(32, e - k0);
            for (int j = 0; j < m; ++j) sum = __dadd_rn(sum, __shfl_sync(0xffffffffu, p, j));
          }
          if (lane == 0) spmv_row_epilogue<MODE>(a, row, sum, spmv_prefetch_aux<MODE>(a, row), dot);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(s_empty + slot);  // this warp is done reading the slot
      PROBE_T(ct2);
      if (tid == 0) {
        PROBE_ADD(1, ct2 - ct1);
        PROBE_ADD(3, 1);
      }
    }
  }
  spmv_finish_dot<MODE>(a, dot, red);
}

}  // namespace b200spai
