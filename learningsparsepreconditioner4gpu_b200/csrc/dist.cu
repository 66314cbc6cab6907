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

#include "pcg_kernels.cuh"

namespace b200spai {

constexpr int MAXR = 16;
enum { V_P = 0, V_R = 1, V_T = 2, NVEC = 3 };
enum { RED_INIT = 0, RED_RHO = 1, RED_PQ = 2, RED_RR = 3, NRED = 4 };

// header of the symmetric segment (same offsets on every rank)
struct SegHeader {
  unsigned long long halo_flag[NVEC][MAXR];
  unsigned long long red_epoch[NRED][MAXR];
  double red_val[NRED][MAXR];
};
constexpr size_t SEG_VEC_OFFSET = (sizeof(SegHeader) + 255) / 256 * 256;

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

// Wait until *flag >= epoch.  On timeout the solve is marked broken so that every later kernel returns at once.
__device__ __forceinline__ void dist_wait(const DistCtx& c, const unsigned long long* flag, unsigned long long epoch) {
  const long long t0 = clock64();
  while (ld_acquire_sys(flag) < epoch) {
    if (clock64() - t0 > c.spin_limit) {
      c.sc->breakdown = 3;
      c.sc->done = 1;
      if (c.sc->use_cond) cudaGraphSetConditional((cudaGraphConditionalHandle)c.sc->cond_handle, 0u);
      break;
    }
  }
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
  // every CTA's peer stores precede its ticket; this CTA drew the last one
  __threadfence_system();
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
__global__ void __launch_bounds__(32) dist_reduce_kernel(const __grid_constant__ DistCtx c, int kind) {
  __shared__ double s_val[MAXR];
  __shared__ unsigned long long s_epoch;
  PcgScalars* sc = c.sc;
  if (kind != RED_INIT && sc->done) return;
  const int t = threadIdx.x;
  if (t == 0) s_epoch = ++c.epoch[NVEC + kind];
  __syncwarp();
  const unsigned long long ep = s_epoch;
  const double local = sc->local_total;
  if (t < c.world) {
    c.hdr[t]->red_val[kind][c.rank] = local;
    __threadfence_system();
    st_release_sys(&c.hdr[t]->red_epoch[kind][c.rank], ep);
  }
  if (t < c.world) {
    dist_wait(c, &c.hdr[c.rank]->red_epoch[kind][t], ep);
    s_val[t] = *reinterpret_cast<volatile double*>(&c.hdr[c.rank]->red_val[kind][t]);
  }
  __syncwarp();
  if (t != 0) return;
  if (sc->breakdown == 3) return;  // a wait timed out
  double total = 0.0;
  for (int r = 0; r < c.world; ++r) total += s_val[r];
  const int was_dist = sc->dist;
  sc->dist = 0;  // the shared update code below must apply the update, not store a local total
  if (kind == RED_INIT) pcg_final_init(sc, total, c.hist);
  else if (kind == RED_RHO) pcg_final_scalar(sc, FINAL_RHO, total);
  else if (kind == RED_PQ) pcg_final_scalar(sc, FINAL_PQ, total);
  else pcg_final_rr(sc, total, c.hist);
  sc->dist = was_dist;
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

static int dist_enqueue_reduce(b200spai_dist_s* s, int kind, cudaStream_t st, int* nl) {
  dist_reduce_kernel<<<1, 32, 0, st>>>(s->ctx, kind);
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

static int dist_enqueue_iteration(b200spai_dist_s* s, int precond, double* x, cudaStream_t st, int* nl) {
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
    B2_TRY(dist_enqueue_halo(s, V_T, st, nl));
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
  B2_TRY(dist_enqueue_reduce(s, RED_RHO, st, nl));
  pcg_update_p_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(s->z, p, s->n, s->sc);
  B2_LAUNCH_CHECK();
  *nl += 1;
  B2_TRY(dist_enqueue_halo(s, V_P, st, nl));
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
  B2_TRY(dist_enqueue_reduce(s, RED_PQ, st, nl));
  pcg_update_xr_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(x, r, p, s->q, s->n, s->sc, s->partials, s->ticket, s->hist);
  B2_LAUNCH_CHECK();
  *nl += 1;
  B2_TRY(dist_enqueue_reduce(s, RED_RR, st, nl));
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) B2_TRY(dist_enqueue_halo(s, V_R, st, nl));  // the next t = G^T r gathers r
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

static int dist_build_graph(b200spai_dist_s* s, int precond, cudaStream_t st) {
  if (s->exec[precond]) return 0;
  cudaGraph_t graph = nullptr;
  cudaStream_t cs = s->cap_stream;
  B2_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeRelaxed));
  int nl0 = 0;
  int rc = dist_enqueue_init(s, s->bw, s->xw, cs, &nl0);
  if (rc) {
    cudaGraph_t junk = nullptr;
    cudaStreamEndCapture(cs, &junk);
    if (junk) cudaGraphDestroy(junk);
    return rc;
  }
  cudaStreamCaptureStatus status;
  const cudaGraphNode_t* deps = nullptr;
  size_t ndeps = 0;
  B2_CUDA(cudaStreamGetCaptureInfo(cs, &status, nullptr, &graph, &deps, &ndeps));
  cudaGraphConditionalHandle handle = (cudaGraphConditionalHandle)s->cond[precond];
  cudaGraphNodeParams cp = {};
  cp.type = cudaGraphNodeTypeConditional;
  cp.conditional.handle = handle;
  cp.conditional.type = cudaGraphCondTypeWhile;
  cp.conditional.size = 1;
  cudaGraphNode_t cnode;
  B2_CUDA(cudaGraphAddNode(&cnode, graph, deps, ndeps, &cp));
  cudaGraph_t body = cp.conditional.phGraph_out[0];
  B2_CUDA(cudaStreamUpdateCaptureDependencies(cs, &cnode, 1, cudaStreamSetCaptureDependencies));
  int nl = 0;
  B2_CUDA(cudaStreamBeginCaptureToGraph(s->body_stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
  rc = dist_enqueue_iteration(s, precond, s->xw, s->body_stream, &nl);
  cudaGraph_t body_out = nullptr;
  cudaError_t e1 = cudaStreamEndCapture(s->body_stream, &body_out);
  cudaGraph_t full = nullptr;
  cudaError_t e2 = cudaStreamEndCapture(cs, &full);
  if (rc) return rc;
  B2_CUDA(e1);
  B2_CUDA(e2);
  cudaGraphExec_t exec = nullptr;
  B2_CUDA(cudaGraphInstantiate(&exec, full, 0));
  s->graph[precond] = full;
  s->exec[precond] = exec;
  s->launches_per_iter[precond] = nl;
  return 0;
}

}  // namespace b200spai
