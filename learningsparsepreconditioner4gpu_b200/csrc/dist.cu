// Row-partitioned PCG for ONE large system over the GPUs of an NVSwitch box.
//
// One process per GPU.  Rank r owns the contiguous rows [rs, re) of A, G and G^T (local CSR matrices whose column
// indices are already remapped to the rank's EXTENDED vector layout: global column c lives at c - shift, the
// extended range [shift, hi) covers the owned rows plus every column the three local matrices touch).  The three
// gathered vectors (p, r, t) live in a peer-mapped "symmetric" segment (cudaMalloc + CUDA IPC, handles exchanged by
// the caller over torch.distributed); everything below is device-driven, no NCCL call and no host involvement
// inside the loop:
//
//  * halo exchange = dist_halo_kernel: every rank STORES the owned entries its neighbours need straight into their
//    extended vectors over NVLink (peer stores), fences, the last CTA raises an epoch flag on each neighbour
//    (st.release.sys) and then waits for the flags of the ranks it receives from (ld.acquire.sys) -- so the SpMV that
//    follows in stream order finds its halo in place;
//  * all-reduce of a CG scalar = dist_reduce_kernel: each rank stores its local total into slot [rank] of every
//    peer, raises the epoch, waits for all slots and sums them IN RANK ORDER -- every rank obtains the same bits, so
//    alpha, beta, the iteration count and the convergence flag are identical everywhere and the ranks leave the
//    loop together without talking to the host.
//
// The five hot kernels are exactly those of the single-GPU solver (pcg_kernels.cuh, spmv.cuh); with sc->dist set
// their last CTA stops at the local total and the all-reduce kernel applies the scalar update.
// Slot/flag reuse is safe because two other all-reduces separate consecutive uses of a slot.
#include <cstdlib>
#include <cstring>
#include <vector>

#include "pcg_fused.cuh"

namespace b200spai {

constexpr int MAXR = 16;
static_assert(MAXR == FUSED_MAXR, "the fused kernel's exchange block is sized for the same number of ranks");
enum { V_P = 0, V_R = 1, V_T = 2, V_Z = 3, V_Q = 4, NVEC = 5 };
enum { RED_INIT = 0, RED_RHO = 1 /* + ||r||^2 */, RED_PQ = 2, NRED = 3 };

// header of the symmetric segment (same offsets on every rank)
struct SegHeader {
  unsigned long long halo_flag[NVEC][MAXR];
  unsigned long long red_epoch[NRED][MAXR];
  double red_val[NRED][MAXR];
  double red_val2[NRED][MAXR];  // second scalar of a fused exchange (||r||^2 travelling with rho)
};
// segment = SegHeader | FusedXchg (exchange block of the fused persistent kernel, mode 4) | NVEC extended vectors
constexpr size_t SEG_XCHG_OFFSET = (sizeof(SegHeader) + 255) / 256 * 256;
constexpr size_t SEG_VEC_OFFSET = SEG_XCHG_OFFSET + (sizeof(FusedXchg) + 255) / 256 * 256;

struct DistCtx {
  int rank, world, n_push, n_recv;
  int push_peer[MAXR];
  long long push_dst_off[MAXR], push_src_off[MAXR];
  int push_len[MAXR];
  int recv_peer[MAXR];
  double* vec[NVEC][MAXR];  // extended vectors of every rank (own included)
  SegHeader* hdr[MAXR];     // segment headers of every rank (own included)
  unsigned long long* epoch;  // [NVEC + NRED] local counters
  unsigned* ticket;           // local, for the halo kernel
  PcgScalars* sc;
  double* hist;
  long long spin_limit;  // clock64 ticks before a wait gives up (a peer died): ~10 s
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }

// Wait until *flag >= epoch: relaxed polls (a system-scope acquire costs microseconds, so it is not repeated per
// poll) and ONE acquire fence once the flag is there.  On timeout the solve is marked broken so that every later
// kernel returns at once.
// Returns false on timeout.
__device__ __forceinline__ bool dist_wait(const DistCtx& c, const unsigned long long* flag, unsigned long long epoch) {
  const long long t0 = clock64();
  bool ok = true;
  while (ld_relaxed_sys(flag) < epoch) {
    if (clock64() - t0 > c.spin_limit) {
      c.sc->breakdown = 3;
      c.sc->done = 1;
      if (c.sc->use_cond) cudaGraphSetConditional((cudaGraphConditionalHandle)c.sc->cond_handle, 0u);
      ok = false;
      break;
    }
  }
  fence_acq_rel_sys();
  return ok;
}

// ---- halo exchange of one extended vector ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) dist_halo_kernel(const __grid_constant__ DistCtx c, int v) {
  __shared__ int s_last;
  __shared__ unsigned long long s_epoch;
  if (c.sc->done) return;
  const double* mine = c.vec[v][c.rank];
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (int s = 0; s < c.n_push; ++s) {
    double* __restrict__ dst = c.vec[v][c.push_peer[s]] + c.push_dst_off[s];
    const double* __restrict__ src = mine + c.push_src_off[s];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < c.push_len[s]; i += stride) dst[i] = src[i];
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(c.ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  // every CTA's peer stores precede (fence.sys) its ticket; this CTA drew the last one, and the release stores below
  // (fence.sys + store) are cumulative over what the ticket atomics made visible to it
  if (threadIdx.x == 0) {
    *c.ticket = 0u;
    s_epoch = ++c.epoch[v];
  }
  __syncthreads();
  const unsigned long long ep = s_epoch;
  if (threadIdx.x < c.n_push) st_release_sys(&c.hdr[c.push_peer[threadIdx.x]]->halo_flag[v][c.rank], ep);
  if (threadIdx.x < c.n_recv) dist_wait(c, &c.hdr[c.rank]->halo_flag[v][c.recv_peer[threadIdx.x]], ep);
}

// ---- all-reduce (sum, fp64, rank order) of sc->local_total + the scalar update it feeds -----------------------
// RED_RHO also carries ||r||^2 of the preceding x/r update (sc->have_rr): the convergence test is applied first, then
// rho / beta -- the order of the single-GPU solver -- so the t = G^T r and z = G t kernels of an iteration that turns out
// not to be needed have run speculatively (they only write t and z) and the exchange count drops from 3 to 2.
// With v >= 0 the same launch first pushes the owned boundary slices of extended vector v into the neighbours (all
// CTAs, like dist_halo_kernel); the CTA that draws the last ticket then runs the all-reduce, whose release / acquire
// flags order the pushed slices as well -- z travels with rho and q with p.q, so an iteration has 3 synchronisation
// points (t, rho + z, p.q + q) instead of 5; p and r are updated on the halo by the vector kernels themselves.
__global__ void __launch_bounds__(256) dist_reduce_kernel(const __grid_constant__ DistCtx c, int kind, int v) {
  __shared__ double s_val[MAXR], s_val2[MAXR];
  __shared__ unsigned long long s_epoch;
  __shared__ int s_last;
  PcgScalars* sc = c.sc;
  if (kind != RED_INIT && sc->done) return;
  if (v >= 0 && c.n_push > 0) {
    const double* mine = c.vec[v][c.rank];
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (int s = 0; s < c.n_push; ++s) {
      double* __restrict__ dst = c.vec[v][c.push_peer[s]] + c.push_dst_off[s];
      const double* __restrict__ src = mine + c.push_src_off[s];
      for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < c.push_len[s]; i += stride) dst[i] = src[i];
    }
    __threadfence_system();
  }
  if (gridDim.x > 1) {
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned tk = atomicAdd(c.ticket, 1u);
      s_last = (tk == gridDim.x - 1);
      if (s_last) *c.ticket = 0u;
    }
    __syncthreads();
    if (!s_last) return;  // every CTA's peer stores precede (fence.sys) its ticket; this CTA drew the last one
  }
  if (threadIdx.x >= 32) return;
  const int t = threadIdx.x;
  // Lane 0 takes a private copy of the scalar block NOW (independent loads, in flight while the exchange runs): the
  // scalar update after the exchange then works on registers and is written back with independent stores, instead of
  // a chain of dependent L2 round trips on the critical path of every iteration.
  PcgScalars loc;
  if (t == 0) {
    loc = *sc;
    s_epoch = ++c.epoch[NVEC + kind];
  }
  __syncwarp();
  const unsigned long long ep = s_epoch;
  const double local = sc->local_total;
  const bool with_rr = (kind == RED_RHO) && sc->have_rr;  // identical on every rank
  const double local2 = with_rr ? sc->local_rr : 0.0;
  bool ok = true;
  if (t < c.world) {
    c.hdr[t]->red_val[kind][c.rank] = local;
    if (with_rr) c.hdr[t]->red_val2[kind][c.rank] = local2;
    st_release_sys(&c.hdr[t]->red_epoch[kind][c.rank], ep);  // orders the two stores above (same thread) and the pushes
    ok = dist_wait(c, &c.hdr[c.rank]->red_epoch[kind][t], ep);
    s_val[t] = *reinterpret_cast<volatile double*>(&c.hdr[c.rank]->red_val[kind][t]);
    if (with_rr) s_val2[t] = *reinterpret_cast<volatile double*>(&c.hdr[c.rank]->red_val2[kind][t]);
  }
  __syncwarp();
  ok = __all_sync(0xffffffffu, ok);
  if (t != 0 || !ok) return;  // a wait timed out: the solve is already marked broken in global memory
  double total = 0.0, total2 = 0.0;
  for (int r = 0; r < c.world; ++r) total += s_val[r];
  if (with_rr)
    for (int r = 0; r < c.world; ++r) total2 += s_val2[r];
  loc.dist = 0;  // the shared update code below must apply the update, not store a local total
  if (kind == RED_INIT) {
    pcg_final_init(&loc, total, c.hist);
  } else if (kind == RED_RHO) {
    if (with_rr) {
      loc.have_rr = 0;
      pcg_final_rr(&loc, total2, c.hist);
    }
    if (!loc.done) pcg_final_scalar(&loc, FINAL_RHO, total);
  } else {
    pcg_final_scalar(&loc, FINAL_PQ, total);
  }
  loc.dist = 1;
  *sc = loc;
}

// diag[i] = A_local[i, i + off]
__global__ void dist_diagonal_kernel(CsrView A, int off, double* __restrict__ diag) {
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= A.n_rows) return;
  double d = 0.0;
  for (int k = A.rowptr[row]; k < A.rowptr[row + 1]; ++k)
    if (A.col[k] == row + off) d += A.val[k];
  diag[row] = d;
}

// rows [r0, r1) of a CSR matrix with every column shifted by -shift
__global__ void slice_rowptr_kernel(const int* __restrict__ rowptr, int r0, int n_local, int* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= n_local) out[i] = rowptr[r0 + i] - rowptr[r0];
}
__global__ void slice_entries_kernel(const int* __restrict__ col, const double* __restrict__ val, long long base, long long cnt,
                                     int shift, int n_cols_local, int* __restrict__ col_out, double* __restrict__ val_out,
                                     int* __restrict__ bad) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < cnt; k += stride) {
    const int c = col[base + k] - shift;
    if (c < 0 || c >= n_cols_local) atomicOr(bad, 1);
    col_out[k] = c;
    val_out[k] = val[base + k];
  }
}

}  // namespace b200spai

using namespace b200spai;

struct b200spai_dist_s {
  int rank = 0, world = 1;
  long long ext_cap = 0;   // capacity (doubles) of each extended vector, identical on every rank
  unsigned char* seg = nullptr;  // local symmetric segment
  size_t seg_bytes = 0;
  void* peer_seg[MAXR] = {nullptr};  // mapped segments (own = seg)
  bool connected = false;
  long long rs = 0, re = 0, shift = 0, hi = 0;  // owned rows, extended range [shift, hi)
  DistCtx ctx{};
  // solver
  b200spai_csr_s *A = nullptr, *G = nullptr, *GT = nullptr;
  double eps = 0.0;
  int n = 0, off = 0;  // owned rows, offset of the owned slice inside the extended vectors
  double *z = nullptr, *q = nullptr, *xw = nullptr, *bw = nullptr, *d = nullptr, *partials = nullptr, *hist = nullptr;
  bool have_d = false;
  size_t vec_cap = 0;  // capacity (doubles) of xw, bw
  long long hist_cap = 0;
  PcgScalars* sc = nullptr;
  unsigned *ticket = nullptr, *ticket_halo = nullptr;
  unsigned long long* epoch = nullptr;
  PcgScalars *h_in = nullptr, *h_out = nullptr;
  int* h_done = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_chunk[2] = {nullptr, nullptr};
  cudaStream_t body_stream = nullptr, cap_stream = nullptr;
  cudaGraphExec_t exec[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaGraph_t graph[4] = {nullptr, nullptr, nullptr, nullptr};
  unsigned long long cond[4] = {0, 0, 0, 0};
  int launches_per_iter[4] = {0, 0, 0, 0};
  int mode = 0;
  long long launches = 0;
  // fused persistent kernel (mode 4): barrier block, phase timers, epoch counters, plan
  FusedSync* fsync = nullptr;
  unsigned long long *fprof = nullptr, *fepoch = nullptr;
  FusedPlan fplan;
  bool fplan_valid = false;
  int4* fblk[3] = {nullptr, nullptr, nullptr};  // row blocks of A, G, G^T with the boundary blocks first (mode 4)
  int fnbnd[3] = {0, 0, 0};                     // how many of them are boundary blocks
  double* vec(int v) const { return reinterpret_cast<double*>(seg + SEG_VEC_OFFSET) + (size_t)v * (size_t)ext_cap; }
};

namespace b200spai {

static void dist_drop_graphs(b200spai_dist_s* s) {
  for (int i = 0; i < 4; ++i) {
    if (s->exec[i]) cudaGraphExecDestroy(s->exec[i]);
    if (s->graph[i]) cudaGraphDestroy(s->graph[i]);
    s->exec[i] = nullptr;
    s->graph[i] = nullptr;
  }
}

static int halo_grid(const b200spai_dist_s* s) {
  long long total = 0;
  for (int i = 0; i < s->ctx.n_push; ++i) total += s->ctx.push_len[i];
  long long g = (total + 256 * 4 - 1) / (256 * 4);
  if (g > 64) g = 64;  // NVLink stores: a few dozen CTAs saturate the two neighbour links of a halo
  if (g < 1) g = 1;
  return (int)g;
}

static int dist_enqueue_halo(b200spai_dist_s* s, int v, cudaStream_t st, int* nl) {
  if (s->world == 1) return 0;
  dist_halo_kernel<<<halo_grid(s), 256, 0, st>>>(s->ctx, v);
  B2_LAUNCH_CHECK();
  *nl += 1;
  return 0;
}

static int dist_enqueue_reduce(b200spai_dist_s* s, int kind, cudaStream_t st, int* nl, int v = -1) {
  const bool push = (v >= 0 && s->world > 1 && s->ctx.n_push > 0);
  dist_reduce_kernel<<<push ? halo_grid(s) : 1, 256, 0, st>>>(s->ctx, kind, push ? v : -1);
  B2_LAUNCH_CHECK();
  *nl += 1;
  return 0;
}

static int dist_enqueue_init(b200spai_dist_s* s, const double* b, double* x, cudaStream_t st, int* nl) {
  double* r = s->vec(V_R) + s->off;
  pcg_init_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(b, x, r, s->n, s->sc, s->partials, s->ticket, s->hist);
  B2_LAUNCH_CHECK();
  *nl += 1;
  B2_TRY(dist_enqueue_reduce(s, RED_INIT, st, nl));
  B2_TRY(dist_enqueue_halo(s, V_R, st, nl));
  return 0;
}

// marks: optional array of 8 events recorded after the launches of one iteration, in the order
// {t = G^T r, halo t, z = G t (+ r.z), all-reduce rho + push z, p update, q = A p (+ p.q), all-reduce p.q + push q, x/r update}
// (measurement hook b200spai_dist_pcg_profile_iterations).
static int dist_enqueue_iteration(b200spai_dist_s* s, int precond, double* x, cudaStream_t st, int* nl,
                                  cudaEvent_t* marks = nullptr) {
  int mk = 0;
  auto mark = [&]() {
    if (marks) cudaEventRecord(marks[mk], st);
    ++mk;
  };
  double *r_ext = s->vec(V_R), *p_ext = s->vec(V_P), *t_ext = s->vec(V_T);
  double *r = r_ext + s->off, *p = p_ext + s->off, *t = t_ext + s->off;
  if (precond == B200SPAI_PRECOND_NONE || precond == B200SPAI_PRECOND_DIAGONAL) {
    pcg_jacobi_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(r, precond == B200SPAI_PRECOND_DIAGONAL ? s->d : nullptr, s->z, s->n,
                                                           s->sc, s->partials, s->ticket);
    B2_LAUNCH_CHECK();
    *nl += 1;
  } else {
    const bool scaled = (precond == B200SPAI_PRECOND_EXT_SPAI_SCALED);
    SpmvArgs a1{};
    a1.x = r_ext;
    a1.y = t;
    a1.d = scaled ? s->d : nullptr;
    a1.sc = s->sc;
    B2_TRY(spmv_launch(s->GT, a1, scaled ? SPMV_SCALE : SPMV_PLAIN, 0, st));
    *nl += 1;
    mark();
    B2_TRY(dist_enqueue_halo(s, V_T, st, nl));
    mark();
    SpmvArgs a2{};
    a2.x = t_ext;
    a2.y = s->z;
    a2.aux = r;
    a2.d = scaled ? s->d : nullptr;
    a2.eps = s->eps;
    a2.partials = s->partials;
    a2.ticket = s->ticket;
    a2.sc = s->sc;
    a2.final_op = FINAL_RHO;
    B2_TRY(spmv_launch(s->G, a2, SPMV_AXPY_DOT, 0, st));
    *nl += 1;
  }
  mk = 2;
  mark();
  const int lo_halo = s->off, hi_halo = (int)(s->hi - s->re);
  B2_TRY(dist_enqueue_reduce(s, RED_RHO, st, nl, V_Z));  // z's boundary travels with rho
  mark();
  pcg_update_p_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(s->z, p, s->n, s->sc, lo_halo, hi_halo);
  B2_LAUNCH_CHECK();
  *nl += 1;
  mark();
  SpmvArgs a{};
  a.x = p_ext;
  a.aux = p;
  a.y = s->q;
  a.partials = s->partials;
  a.ticket = s->ticket;
  a.sc = s->sc;
  a.final_op = FINAL_PQ;
  B2_TRY(spmv_launch(s->A, a, SPMV_DOT, 0, st));
  *nl += 1;
  mark();
  // q's boundary travels with p.q (only the SPAI preconditioners gather r, through t = G^T r);
  // ||r||^2 is all-reduced together with the next rho (dist_reduce_kernel)
  const bool need_r = precond >= B200SPAI_PRECOND_EXT_SPAI;
  B2_TRY(dist_enqueue_reduce(s, RED_PQ, st, nl, need_r ? V_Q : -1));
  mark();
  pcg_update_xr_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(x, r, p, s->q, s->n, s->sc, s->partials, s->ticket, s->hist,
                                                            need_r ? lo_halo : 0, need_r ? hi_halo : 0);
  B2_LAUNCH_CHECK();
  *nl += 1;
  mark();
  return 0;
}

static int dist_ensure_hist(b200spai_dist_s* s, long long need) {
  if (need <= s->hist_cap) return 0;
  cudaFree(s->hist);
  s->hist = nullptr;
  B2_TRY(dev_alloc(&s->hist, (size_t)need));
  s->hist_cap = need;
  s->ctx.hist = s->hist;
  dist_drop_graphs(s);  // graphs bake the context (kernel parameter)
  return 0;
}

static int dist_warm(b200spai_dist_s* s, int precond, cudaStream_t st) {
  // build the lazily created row partitions / compressed streams outside any capture (their build synchronises)
  SpmvArgs w{};
  w.x = s->vec(V_P);
  w.y = s->q;
  B2_TRY(spmv_launch(s->A, w, SPMV_PLAIN, 0, st));
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) {
    w.x = s->vec(V_R);
    w.y = s->z;
    B2_TRY(spmv_launch(s->GT, w, SPMV_PLAIN, 0, st));
    B2_TRY(spmv_launch(s->G, w, SPMV_PLAIN, 0, st));
  }
  B2_CUDA(cudaStreamSynchronize(st));
  return 0;
}

}  // namespace b200spai

// =================================================================================================================
// C ABI
// =================================================================================================================
extern "C" {

int b200spai_dist_create(int32_t rank, int32_t world, int64_t ext_cap, b200spai_dist_t* out) {
  B2_ARG(out && world >= 1 && world <= MAXR && rank >= 0 && rank < world && ext_cap > 0, "dist_create: bad argument");
  auto* s = new b200spai_dist_s();
  s->rank = rank;
  s->world = world;
  s->ext_cap = (ext_cap + 31) & ~31LL;
  s->seg_bytes = SEG_VEC_OFFSET + (size_t)NVEC * (size_t)s->ext_cap * sizeof(double);
  // plain cudaMalloc (not the pool): the segment is exported to the peer processes through CUDA IPC
  cudaError_t e = cudaMalloc((void**)&s->seg, s->seg_bytes);
  if (e == cudaSuccess) e = cudaMemset(s->seg, 0, s->seg_bytes);
  if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamLegacy);
  if (e != cudaSuccess) {
    set_error(std::string("dist_create: ") + cudaGetErrorString(e));
    delete s;
    return B200SPAI_ERR_CUDA;
  }
  s->peer_seg[rank] = s->seg;
  *out = s;
  return 0;
}

int b200spai_dist_ipc_handle(b200spai_dist_t s, unsigned char* handle64) {
  B2_ARG(s && handle64, "dist_ipc_handle: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
  cudaIpcMemHandle_t h;
  B2_CUDA(cudaIpcGetMemHandle(&h, s->seg));
  memcpy(handle64, &h, 64);
  return 0;
}

// handles: [world][64] as gathered from every rank; layout: [world][4] = {row_start, row_end, shift, hi} per rank
// (extended range [shift, hi); shift may be row-aligned down by one so that the owned slice starts 16-byte aligned).
// The layout-dependent part of the connection: owned rows, extended range, push / receive plan, local vectors.
static int dist_apply_layout(b200spai_dist_s* s, const int64_t* L) {
  const int R = s->world, me = s->rank;
  const int64_t rs = L[me * 4 + 0], re = L[me * 4 + 1], shift = L[me * 4 + 2], hi = L[me * 4 + 3];
  B2_ARG(rs <= re && shift <= rs && hi >= re, "dist layout: inconsistent layout");
  B2_ARG(hi - shift <= s->ext_cap, "dist layout: extended range exceeds the segment capacity");
  B2_ARG(((rs - shift) & 1) == 0, "dist layout: the owned slice must start at an even extended index");
  s->rs = rs;
  s->re = re;
  s->shift = shift;
  s->hi = hi;
  s->n = (int)(re - rs);
  s->off = (int)(rs - shift);
  DistCtx& c = s->ctx;
  c.rank = me;
  c.world = R;
  c.n_push = c.n_recv = 0;
  for (int q = 0; q < R; ++q) {
    auto* base = reinterpret_cast<unsigned char*>(s->peer_seg[q]);
    c.hdr[q] = reinterpret_cast<SegHeader*>(base);
    for (int v = 0; v < NVEC; ++v)
      c.vec[v][q] = reinterpret_cast<double*>(base + SEG_VEC_OFFSET) + (size_t)v * (size_t)s->ext_cap;
  }
  for (int q = 0; q < R; ++q) {
    if (q == me) continue;
    const int64_t rs_q = L[q * 4 + 0], re_q = L[q * 4 + 1], lo_q = L[q * 4 + 2], hi_q = L[q * 4 + 3];
    const int64_t a = std::max<int64_t>(s->rs, lo_q), b = std::min<int64_t>(s->re, hi_q);  // what q needs from my rows
    if (a < b) {
      const int k = c.n_push++;
      c.push_peer[k] = q;
      c.push_dst_off[k] = a - lo_q;
      c.push_src_off[k] = a - s->shift;
      c.push_len[k] = (int)(b - a);
    }
    const int64_t a2 = std::max<int64_t>(rs_q, s->shift), b2 = std::min<int64_t>(re_q, s->hi);  // what I need from q
    if (a2 < b2) c.recv_peer[c.n_recv++] = q;
  }
  const size_t n = (size_t)(s->n > 0 ? s->n : 1);
  if (n + 2 > s->vec_cap) {  // local vectors (owned rows)
    int rc = 0;
    for (double** v : {&s->xw, &s->bw}) {
      dev_free(*v);
      *v = nullptr;
      if (!rc) rc = dev_alloc(v, n + 2);
    }
    if (rc) return rc;
    s->vec_cap = n + 2;
  }
  s->z = s->vec(V_Z) + s->off;  // z and q are extended vectors too: their boundaries travel with the all-reduces
  s->q = s->vec(V_Q) + s->off;
  dev_free(s->d);
  s->d = nullptr;
  s->have_d = false;
  s->A = s->G = s->GT = nullptr;
  s->fplan_valid = false;
  dist_drop_graphs(s);
  return 0;
}

int b200spai_dist_connect(b200spai_dist_t s, const unsigned char* handles, const int64_t* layout) {
  B2_ARG(s && layout && (handles || s->world == 1), "dist_connect: NULL argument");
  B2_ARG(!s->connected, "dist_connect: already connected");
  const int R = s->world, me = s->rank;
  for (int q = 0; q < R; ++q) {
    if (q == me) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + (size_t)q * 64, 64);
    B2_CUDA(cudaIpcOpenMemHandle(&s->peer_seg[q], h, cudaIpcMemLazyEnablePeerAccess));
  }
  s->ctx = DistCtx{};
  B2_TRY(dist_apply_layout(s, layout));
  DistCtx& c = s->ctx;
  int rc = dev_alloc(&s->sc, 1);
  if (!rc) rc = dev_alloc(&s->partials, 8192);
  if (!rc) rc = dev_alloc(&s->ticket, 1);
  if (!rc) rc = dev_alloc(&s->ticket_halo, 1);
  if (!rc) rc = dev_alloc(&s->epoch, NVEC + NRED);
  if (rc) return rc;
  B2_CUDA(cudaMemset(s->ticket, 0, sizeof(unsigned)));
  B2_CUDA(cudaMemset(s->ticket_halo, 0, sizeof(unsigned)));
  B2_CUDA(cudaMemset(s->epoch, 0, (NVEC + NRED) * sizeof(unsigned long long)));
  B2_CUDA(cudaMemset(s->sc, 0, sizeof(PcgScalars)));
  B2_CUDA(cudaStreamSynchronize(cudaStreamLegacy));  // default-stream memsets do not order with non-blocking streams
  B2_CUDA(cudaMallocHost((void**)&s->h_in, sizeof(PcgScalars)));
  B2_CUDA(cudaMallocHost((void**)&s->h_out, sizeof(PcgScalars)));
  B2_CUDA(cudaMallocHost((void**)&s->h_done, 2 * sizeof(int)));
  B2_CUDA(cudaEventCreate(&s->ev0));
  B2_CUDA(cudaEventCreate(&s->ev1));
  B2_CUDA(cudaEventCreateWithFlags(&s->ev_chunk[0], cudaEventDisableTiming));
  B2_CUDA(cudaEventCreateWithFlags(&s->ev_chunk[1], cudaEventDisableTiming));
  B2_CUDA(cudaStreamCreateWithFlags(&s->body_stream, cudaStreamNonBlocking));
  B2_CUDA(cudaStreamCreateWithFlags(&s->cap_stream, cudaStreamNonBlocking));
  c.epoch = s->epoch;
  c.ticket = s->ticket_halo;
  c.sc = s->sc;
  c.hist = nullptr;
  int khz = 1965000;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const char* e = getenv("B200SPAI_DIST_TIMEOUT_S");
  const double secs = e ? atof(e) : 10.0;
  c.spin_limit = (long long)(secs * 1e3 * (double)khz);
  s->connected = true;
  return 0;
}

// A connected handle is reusable for the next system (same ranks, another partition): the new layout of every rank,
// collective like dist_connect.  The mapped segments, the epoch counters and the streams are kept, so the per-system
// cost is the index arithmetic of the plan; the matrices must be attached again with b200spai_dist_pcg_setup.
int b200spai_dist_relayout(b200spai_dist_t s, const int64_t* layout) {
  B2_ARG(s && layout && s->connected, "dist_relayout: not connected or NULL layout");
  B2_CUDA(cudaDeviceSynchronize());  // no kernel of the previous solve may still use the old plan
  // stale vectors of the previous system: a peer only writes into this segment after the first all-reduce of the
  // next solve, which needs this rank's contribution, so clearing here cannot race with its pushes
  B2_CUDA(cudaMemset(s->seg + SEG_VEC_OFFSET, 0, s->seg_bytes - SEG_VEC_OFFSET));
  B2_CUDA(cudaStreamSynchronize(cudaStreamLegacy));
  return dist_apply_layout(s, layout);
}

int b200spai_dist_info(b200spai_dist_t s, int32_t* n_push, int32_t* n_recv, int64_t* push_elems, int32_t* owned_offset) {
  B2_ARG(s && s->connected, "dist_info: not connected");
  if (n_push) *n_push = s->ctx.n_push;
  if (n_recv) *n_recv = s->ctx.n_recv;
  if (push_elems) {
    int64_t t = 0;
    for (int i = 0; i < s->ctx.n_push; ++i) t += s->ctx.push_len[i];
    *push_elems = t;
  }
  if (owned_offset) *owned_offset = s->off;
  return 0;
}

/* Rows [r0, r1) of `a` as a new matrix with n_cols_local columns, every column index shifted by -shift. */
int b200spai_csr_slice_rows(b200spai_csr_t a, int32_t r0, int32_t r1, int64_t shift, int32_t n_cols_local, void* stream,
                            b200spai_csr_t* out) {
  B2_ARG(a && out && r0 >= 0 && r0 <= r1 && r1 <= a->n_rows && n_cols_local >= 0, "csr_slice_rows: bad argument");
  cudaStream_t st = as_stream(stream);
  int h[2] = {0, 0};
  B2_CUDA(cudaMemcpyAsync(&h[0], a->rowptr + r0, sizeof(int), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaMemcpyAsync(&h[1], a->rowptr + r1, sizeof(int), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  const long long cnt = (long long)h[1] - h[0];
  b200spai_csr_s* L = nullptr;
  B2_TRY(csr_alloc(r1 - r0, n_cols_local, cnt, &L));
  if (a->block_hint == 3 && r0 % 3 == 0 && (r1 - r0) % 3 == 0 && shift % 3 == 0 && n_cols_local % 3 == 0) L->block_hint = 3;
  int* bad = nullptr;
  B2_TRY(dev_alloc(&bad, 1));
  B2_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), st));
  const int nl = r1 - r0;
  slice_rowptr_kernel<<<(nl + 1 + 255) / 256, 256, 0, st>>>(a->rowptr, r0, nl, L->rowptr);
  if (cnt > 0) {
    long long g = (cnt + 1023) / 1024;
    if (g > (long long)sm_count() * 8) g = (long long)sm_count() * 8;
    slice_entries_kernel<<<(int)g, 256, 0, st>>>(a->col, a->val, h[0], cnt, (int)shift, n_cols_local, L->col, L->val, bad);
  }
  B2_LAUNCH_CHECK();
  count_launches(cnt > 0 ? 2 : 1);
  int hb = 0;
  B2_CUDA(cudaMemcpyAsync(&hb, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  dev_free(bad);
  if (hb) {
    csr_free(L);
    set_error("csr_slice_rows: a column falls outside [shift, shift + n_cols_local)");
    return B200SPAI_ERR_ARG;
  }
  *out = L;
  return 0;
}

int b200spai_dist_pcg_setup(b200spai_dist_t s, b200spai_csr_t a_local, b200spai_csr_t g_local, b200spai_csr_t gt_local,
                            double epsilon) {
  B2_ARG(s && s->connected && a_local, "dist_pcg_setup: not connected or NULL matrix");
  const int ext = (int)(s->hi - s->shift);
  B2_ARG(a_local->n_rows == s->n && a_local->n_cols == ext, "dist_pcg_setup: A_local must be [owned rows x extended range]");
  if (g_local || gt_local) {
    B2_ARG(g_local && gt_local, "dist_pcg_setup: G and G^T come together");
    B2_ARG(g_local->n_rows == s->n && g_local->n_cols == ext && gt_local->n_rows == s->n && gt_local->n_cols == ext,
           "dist_pcg_setup: G_local / GT_local must be [owned rows x extended range]");
  }
  s->A = a_local;
  s->G = g_local;
  s->GT = gt_local;
  s->eps = epsilon;
  s->have_d = false;
  s->fplan_valid = false;
  dist_drop_graphs(s);
  return 0;
}

int b200spai_dist_pcg_set_mode(b200spai_dist_t s, int mode) {
  B2_ARG(s && (mode == 0 || mode == 1 || mode == 2 || mode == 4), "dist_pcg_set_mode: mode must be 0, 1, 2 or 4");
  s->mode = mode;
  return 0;
}

// The row blocks whose rows a neighbour needs go FIRST in every CTA's sweep: their peer stores are long complete when the
// sweep ends, so the system-scope fence before the barrier finds nothing in flight (in row order the high boundary would
// be written last, right before the barrier).
static int dist_boundary_first(b200spai_dist_s* s, const b200spai_csr_s* M, int which, int4** out, int* n_boundary, cudaStream_t st) {
  const int nb = M->nblk[which];
  dev_free(*out);
  *out = nullptr;
  *n_boundary = 0;
  if (nb <= 0) return 0;
  std::vector<int4> h((size_t)nb), r;
  B2_CUDA(cudaMemcpyAsync(h.data(), M->blk[which], (size_t)nb * sizeof(int4), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  r.reserve((size_t)nb);
  const DistCtx& c = s->ctx;
  auto boundary = [&](const int4& b) {
    for (int k = 0; k < c.n_push; ++k) {
      const long long lo = c.push_src_off[k] - s->off, hi = lo + c.push_len[k];
      if (b.x < hi && b.z > lo) return true;
    }
    return false;
  };
  for (const int4& b : h)
    if (boundary(b)) r.push_back(b);
  *n_boundary = (int)r.size();
  for (const int4& b : h)
    if (!boundary(b)) r.push_back(b);
  B2_TRY(dev_alloc(out, (size_t)nb));
  B2_CUDA(cudaMemcpyAsync(*out, r.data(), (size_t)nb * sizeof(int4), cudaMemcpyHostToDevice, st));
  B2_CUDA(cudaStreamSynchronize(st));  // r is a host temporary
  return 0;
}

static int dist_fused_plan(b200spai_dist_s* s, cudaStream_t st) {
  if (!s->fplan_valid) {
    s->fplan = FusedPlan{};
    if (s->A && s->G && s->GT && s->n > 0) B2_TRY(fused_plan(s->A, s->G, s->GT, st, &s->fplan));
    if (s->fplan.which >= 0 && s->world > 1) {  // (the neighbour flags of barrier B1 rely on this order)
      B2_TRY(dist_boundary_first(s, s->A, s->fplan.which, &s->fblk[0], &s->fnbnd[0], st));
      B2_TRY(dist_boundary_first(s, s->G, s->fplan.which, &s->fblk[1], &s->fnbnd[1], st));
      B2_TRY(dist_boundary_first(s, s->GT, s->fplan.which, &s->fblk[2], &s->fnbnd[2], st));
    }
    s->fplan_valid = true;
  }
  return 0;
}

int b200spai_dist_pcg_fused_ok(b200spai_dist_t s, void* stream, int32_t* ok) {
  B2_ARG(s && s->connected && ok, "dist_pcg_fused_ok: not connected or NULL argument");
  *ok = 0;
  if (!s->A) return 0;
  B2_TRY(dist_fused_plan(s, as_stream(stream)));
  // 2: the matrices are in the streaming regime (nnz >= 2^18), where the automatic choice takes the fused kernel
  *ok = s->fplan.which >= 0 ? ((spmv_auto_kernel(s->A) == 2) ? 2 : 1) : 0;
  return 0;
}

int b200spai_dist_pcg_fused_profile(b200spai_dist_t s, double* phase_ms5, int32_t* iterations, void* stream) {
  B2_ARG(s && phase_ms5, "dist_pcg_fused_profile: NULL argument");
  B2_ARG(s->fprof, "dist_pcg_fused_profile: no fused solve has run on this handle");
  unsigned long long h[6] = {0, 0, 0, 0, 0, 0};
  B2_CUDA(cudaMemcpyAsync(h, s->fprof, sizeof(h), cudaMemcpyDeviceToHost, as_stream(stream)));
  B2_CUDA(cudaStreamSynchronize(as_stream(stream)));
  for (int k = 0; k < 5; ++k) phase_ms5[k] = (double)h[k] * 1e-6;
  if (iterations) *iterations = (int32_t)h[5];
  return 0;
}

int b200spai_dist_pcg_fused_barrier_profile(b200spai_dist_t s, double* wait_us2, int64_t* barriers, void* stream) {
  B2_ARG(s && wait_us2 && barriers, "dist_pcg_fused_barrier_profile: NULL argument");
  B2_ARG(s->fprof, "dist_pcg_fused_barrier_profile: no fused solve has run on this handle");
  unsigned long long h[10] = {0};
  B2_CUDA(cudaMemcpyAsync(h, s->fprof, sizeof(h), cudaMemcpyDeviceToHost, as_stream(stream)));
  B2_CUDA(cudaStreamSynchronize(as_stream(stream)));
  const double nb = h[8] > 0 ? (double)h[8] : 1.0;
  wait_us2[0] = (double)h[6] * 1e-3 / nb;
  wait_us2[1] = (double)h[7] * 1e-3 / nb;
  *barriers = (int64_t)h[8];
  return 0;
}

// Mode 4: the whole row-partitioned solve in ONE persistent kernel per rank (pcg_fused.cuh).
static int dist_solve_fused(b200spai_dist_s* s, int precond, const double* b, double* x, cudaStream_t st) {
  if (!s->fsync) B2_TRY(dev_alloc(&s->fsync, 1));
  if (!s->fprof) B2_TRY(dev_alloc(&s->fprof, 10));
  if (!s->fepoch) {
    B2_TRY(dev_alloc(&s->fepoch, FB_KINDS));
    B2_CUDA(cudaMemsetAsync(s->fepoch, 0, FB_KINDS * sizeof(unsigned long long), st));
  }
  B2_CUDA(cudaMemsetAsync(s->fsync, 0, sizeof(FusedSync), st));
  B2_CUDA(cudaMemsetAsync(s->fprof, 0, 10 * sizeof(unsigned long long), st));
  FusedArgs a{};
  a.A = fused_mat(s->A, s->fplan);
  a.G = fused_mat(s->G, s->fplan);
  a.GT = fused_mat(s->GT, s->fplan);
  if (s->fblk[0]) { a.A.blk = s->fblk[0]; a.A.nbnd = s->fnbnd[0]; }
  if (s->fblk[1]) { a.G.blk = s->fblk[1]; a.G.nbnd = s->fnbnd[1]; }
  if (s->fblk[2]) { a.GT.blk = s->fblk[2]; a.GT.nbnd = s->fnbnd[2]; }
  a.n = s->n;
  a.off = s->off;
  a.hi_halo = (int)(s->hi - s->re);
  a.b = b;
  a.x = x;
  a.r = s->vec(V_R);
  a.p = s->vec(V_P);
  a.z = s->vec(V_Z);
  a.q = s->vec(V_Q);
  a.t = s->vec(V_T);
  a.d = (precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) ? s->d : nullptr;
  a.eps = s->eps;
  a.precond = precond;
  a.sc = s->sc;
  a.partials = s->partials;
  a.hist = s->hist;
  a.sync = s->fsync;
  a.prof = s->fprof;
  a.spin_limit = s->ctx.spin_limit;
  FusedPeer& P = a.peer;
  const DistCtx& c = s->ctx;
  P.rank = s->rank;
  P.world = s->world;
  P.n_push = c.n_push;
  P.n_recv = c.n_recv;
  for (int k = 0; k < c.n_recv; ++k) P.recv_peer[k] = c.recv_peer[k];
  for (int k = 0; k < c.n_push; ++k) {
    const int q = c.push_peer[k];
    P.push_peer[k] = q;
    P.push_lo[k] = (int)(c.push_src_off[k] - s->off);
    P.push_hi[k] = P.push_lo[k] + c.push_len[k];
    P.push_r[k] = c.vec[V_R][q] + c.push_dst_off[k];
    P.push_t[k] = c.vec[V_T][q] + c.push_dst_off[k];
    P.push_z[k] = c.vec[V_Z][q] + c.push_dst_off[k];
    P.push_q[k] = c.vec[V_Q][q] + c.push_dst_off[k];
  }
  for (int q = 0; q < s->world; ++q)
    P.xchg[q] = reinterpret_cast<FusedXchg*>(reinterpret_cast<unsigned char*>(s->peer_seg[q]) + SEG_XCHG_OFFSET);
  P.epoch = s->fepoch;
  B2_TRY(fused_launch(a, s->fplan, st));
  s->launches += 1;
  count_launches(1);
  return 0;
}

int b200spai_dist_pcg_solve(b200spai_dist_t s, int precond, const double* b_local_dev, double* x_local_dev, double rtol,
                            int32_t max_iter, double* resid_hist_dev, void* stream, b200spai_pcg_result* res) {
  B2_ARG(s && s->connected && s->A && b_local_dev && x_local_dev && res && max_iter > 0, "dist_pcg_solve: bad argument");
  B2_ARG(precond >= 0 && precond <= 3, "dist_pcg_solve: unknown preconditioner");
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) B2_ARG(s->G && s->GT, "dist_pcg_solve: ext_spai needs G and G^T");
  cudaStream_t st = as_stream(stream);
  if ((precond == B200SPAI_PRECOND_DIAGONAL || precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) && !s->have_d) {
    if (!s->d) B2_TRY(dev_alloc(&s->d, (size_t)(s->n > 0 ? s->n : 1)));
    if (s->n > 0) dist_diagonal_kernel<<<(s->n + 255) / 256, 256, 0, st>>>(s->A->view(), s->off, s->d);
    B2_LAUNCH_CHECK();
    s->have_d = true;
  }
  B2_TRY(dist_ensure_hist(s, (long long)max_iter + 1));
  int mode = s->mode;
  if (mode == 0) {
    const char* e = getenv("B200SPAI_PCG_MODE");
    mode = (e && atoi(e) == 1) ? 1 : 2;
  }
  if (mode == 4 && precond < B200SPAI_PRECOND_EXT_SPAI) mode = 2;  // the fused kernel is the SPAI iteration; same choice on every rank
  if (mode == 4) {
    B2_TRY(dist_fused_plan(s, st));
    B2_ARG(s->fplan.which >= 0, "dist_pcg_solve: mode 4 does not apply to these matrices (ask b200spai_dist_pcg_fused_ok on every rank)");
  } else {
    B2_TRY(dist_warm(s, precond, st));
  }
  PcgScalars init{};
  init.rtol = rtol;
  init.max_iter = max_iter;
  init.dist = 1;
  if (mode == 2) init.use_cond = 1;
  const size_t nbytes = (size_t)s->n * sizeof(double);
  int nl_init = 0;
  if (mode == 2) {
    if (!s->exec[precond]) {
      // one graph = init (+ its all-reduce and the first r halo) followed by a device-side WHILE over the iteration body
      cudaStream_t cs = s->cap_stream;
      cudaGraph_t graph = nullptr;
      B2_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeRelaxed));
      cudaStreamCaptureStatus status;
      B2_CUDA(cudaStreamGetCaptureInfo(cs, &status, nullptr, &graph, nullptr, nullptr));
      cudaGraphConditionalHandle handle;
      cudaError_t eh = cudaGraphConditionalHandleCreate(&handle, graph, 1, cudaGraphCondAssignDefault);
      if (eh != cudaSuccess) {
        cudaGraph_t junk = nullptr;
        cudaStreamEndCapture(cs, &junk);
        if (junk) cudaGraphDestroy(junk);
        B2_CUDA(eh);
      }
      s->cond[precond] = (unsigned long long)handle;
      // continue the same capture: init + WHILE(body)
      int nl0 = 0;
      int rc = dist_enqueue_init(s, s->bw, s->xw, cs, &nl0);
      const cudaGraphNode_t* deps = nullptr;
      size_t ndeps = 0;
      cudaError_t e0 = cudaStreamGetCaptureInfo(cs, &status, nullptr, &graph, &deps, &ndeps);
      cudaGraphNodeParams cp = {};
      cp.type = cudaGraphNodeTypeConditional;
      cp.conditional.handle = handle;
      cp.conditional.type = cudaGraphCondTypeWhile;
      cp.conditional.size = 1;
      cudaGraphNode_t cnode;
      cudaError_t e1 = (rc == 0 && e0 == cudaSuccess) ? cudaGraphAddNode(&cnode, graph, deps, ndeps, &cp) : cudaErrorUnknown;
      int nl = 0;
      cudaError_t e2 = cudaErrorUnknown, e3 = cudaErrorUnknown;
      if (e1 == cudaSuccess) {
        cudaGraph_t body = cp.conditional.phGraph_out[0];
        e2 = cudaStreamUpdateCaptureDependencies(cs, &cnode, 1, cudaStreamSetCaptureDependencies);
        if (e2 == cudaSuccess) e2 = cudaStreamBeginCaptureToGraph(s->body_stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed);
        if (e2 == cudaSuccess) {
          rc = dist_enqueue_iteration(s, precond, s->xw, s->body_stream, &nl);
          cudaGraph_t body_out = nullptr;
          e3 = cudaStreamEndCapture(s->body_stream, &body_out);
        }
      }
      cudaGraph_t full = nullptr;
      cudaError_t e4 = cudaStreamEndCapture(cs, &full);
      if (rc) return rc;
      B2_CUDA(e0);
      B2_CUDA(e1);
      B2_CUDA(e2);
      B2_CUDA(e3);
      B2_CUDA(e4);
      cudaGraphExec_t exec = nullptr;
      B2_CUDA(cudaGraphInstantiate(&exec, full, 0));
      s->graph[precond] = full;
      s->exec[precond] = exec;
      s->launches_per_iter[precond] = nl;
      nl_init = nl0;
    }
    init.cond_handle = s->cond[precond];
  }
  *s->h_in = init;
  B2_CUDA(cudaEventRecord(s->ev0, st));
  B2_CUDA(cudaMemcpyAsync(s->sc, s->h_in, sizeof(PcgScalars), cudaMemcpyHostToDevice, st));
  if (mode == 4) {
    B2_TRY(dist_solve_fused(s, precond, b_local_dev, x_local_dev, st));
  } else if (mode == 2) {
    if (s->n > 0) B2_CUDA(cudaMemcpyAsync(s->bw, b_local_dev, nbytes, cudaMemcpyDeviceToDevice, st));
    B2_CUDA(cudaGraphLaunch(s->exec[precond], st));
    if (s->n > 0) B2_CUDA(cudaMemcpyAsync(x_local_dev, s->xw, nbytes, cudaMemcpyDeviceToDevice, st));
  } else {
    double* xs = ((reinterpret_cast<uintptr_t>(x_local_dev) & 15) == 0) ? x_local_dev : s->xw;
    int nl = 0;
    B2_TRY(dist_enqueue_init(s, b_local_dev, xs, st, &nl));
    const int CH = 8;
    int slot = 0;
    long long enq = 0;
    bool first = true;
    s->h_done[0] = s->h_done[1] = 0;
    while (true) {
      for (int k = 0; k < CH; ++k) B2_TRY(dist_enqueue_iteration(s, precond, xs, st, &nl));
      enq += CH;
      B2_CUDA(cudaMemcpyAsync(s->h_done + slot, &s->sc->done, sizeof(int), cudaMemcpyDeviceToHost, st));
      B2_CUDA(cudaEventRecord(s->ev_chunk[slot], st));
      slot ^= 1;
      if (first) {
        first = false;
        if (enq < (long long)max_iter + CH) continue;
      }
      // every rank reads the SAME flag values (the scalars are bit-identical), so all ranks stop after the same chunk
      B2_CUDA(cudaEventSynchronize(s->ev_chunk[slot]));
      if (s->h_done[slot]) break;
      if (enq >= (long long)max_iter + 2 * CH) break;
    }
    if (xs != x_local_dev && s->n > 0) B2_CUDA(cudaMemcpyAsync(x_local_dev, xs, nbytes, cudaMemcpyDeviceToDevice, st));
    s->launches += nl;
    count_launches(nl);
  }
  B2_CUDA(cudaEventRecord(s->ev1, st));
  B2_CUDA(cudaMemcpyAsync(s->h_out, s->sc, sizeof(PcgScalars), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  const PcgScalars& o = *s->h_out;
  if (mode == 2) {
    const long long nl = 3 + (long long)s->launches_per_iter[precond] * (o.iter > 0 ? o.iter : 1);
    (void)nl_init;
    s->launches += nl;
    count_launches(nl);
  }
  res->iterations = o.iter;
  res->converged = o.converged;
  res->breakdown = o.breakdown;
  res->reserved = mode;
  res->b_norm = o.bnorm;
  res->final_residual = sqrt(o.rr);
  res->setup_ms = 0.f;
  B2_CUDA(cudaEventElapsedTime(&res->solve_ms, s->ev0, s->ev1));
  if (resid_hist_dev)
    B2_CUDA(cudaMemcpyAsync(resid_hist_dev, s->hist, ((size_t)o.iter + 1) * sizeof(double), cudaMemcpyDefault, st));
  B2_CUDA(cudaStreamSynchronize(st));
  if (o.breakdown == 3) {
    set_error("dist_pcg_solve: a wait on a peer rank timed out (peer process gone?)");
    return B200SPAI_ERR_CUDA;
  }
  return 0;
}

// Measurement hook (collective): `iterations` PCG updates from x0 = 0 with no convergence test, every launch followed
// by a CUDA event; kernel_ms8 = mean device time per launch class in the order of dist_enqueue_iteration's marks
// (a synchronisation kernel's time includes waiting for the slowest neighbour).
int b200spai_dist_pcg_profile_iterations(b200spai_dist_t s, int precond, const double* b_local_dev, double* x_local_dev,
                                         int32_t iterations, float* kernel_ms8, void* stream) {
  B2_ARG(s && s->connected && s->A && b_local_dev && x_local_dev && kernel_ms8 && iterations > 0 && iterations <= 256,
         "dist_pcg_profile_iterations: bad argument");
  B2_ARG(precond >= B200SPAI_PRECOND_EXT_SPAI && s->G && s->GT, "dist_pcg_profile_iterations: needs ext_spai with G and G^T");
  cudaStream_t st = as_stream(stream);
  B2_TRY(dist_ensure_hist(s, (long long)iterations + 2));
  B2_TRY(dist_warm(s, precond, st));
  PcgScalars init{};
  init.rtol = 0.0;
  init.max_iter = iterations + 1;
  init.dist = 1;
  B2_CUDA(cudaMemcpyAsync(s->sc, &init, sizeof(PcgScalars), cudaMemcpyHostToDevice, st));
  std::vector<cudaEvent_t> ev((size_t)iterations * 8 + 1);
  for (auto& e : ev) B2_CUDA(cudaEventCreate(&e));
  int nl = 0;
  int rc = dist_enqueue_init(s, b_local_dev, x_local_dev, st, &nl);
  cudaEventRecord(ev[0], st);
  for (int it = 0; it < iterations && !rc; ++it) rc = dist_enqueue_iteration(s, precond, x_local_dev, st, &nl, &ev[1 + (size_t)it * 8]);
  cudaError_t e = cudaStreamSynchronize(st);
  for (int k = 0; k < 8; ++k) kernel_ms8[k] = 0.0f;
  if (!rc && e == cudaSuccess) {
    for (int it = 0; it < iterations; ++it)
      for (int k = 0; k < 8; ++k) {
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, ev[(size_t)it * 8 + k], ev[(size_t)it * 8 + k + 1]);
        kernel_ms8[k] += ms / (float)iterations;
      }
  }
  for (auto& x : ev) cudaEventDestroy(x);
  s->launches += nl;
  count_launches(nl);
  if (rc) return rc;
  B2_CUDA(e);
  return 0;
}

int b200spai_dist_destroy(b200spai_dist_t s) {
  if (!s) return 0;
  dist_drop_graphs(s);
  for (int q = 0; q < s->world; ++q)
    if (q != s->rank && s->peer_seg[q]) cudaIpcCloseMemHandle(s->peer_seg[q]);
  cudaFree(s->seg);
  for (double* v : {s->xw, s->bw, s->d, s->partials, s->hist}) dev_free(v);
  dev_free(s->sc);
  dev_free(s->fsync);
  dev_free(s->fprof);
  dev_free(s->fepoch);
  for (int4* b : s->fblk) dev_free(b);
  dev_free(s->ticket);
  dev_free(s->ticket_halo);
  dev_free(s->epoch);
  if (s->h_in) cudaFreeHost(s->h_in);
  if (s->h_out) cudaFreeHost(s->h_out);
  if (s->h_done) cudaFreeHost(s->h_done);
  if (s->ev0) cudaEventDestroy(s->ev0);
  if (s->ev1) cudaEventDestroy(s->ev1);
  if (s->ev_chunk[0]) cudaEventDestroy(s->ev_chunk[0]);
  if (s->ev_chunk[1]) cudaEventDestroy(s->ev_chunk[1]);
  if (s->body_stream) cudaStreamDestroy(s->body_stream);
  if (s->cap_stream) cudaStreamDestroy(s->cap_stream);
  delete s;
  return 0;
}

}  // extern "C"
