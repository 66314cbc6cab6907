// Preconditioned conjugate gradients, fp64, with every scalar kept on the device.
//
// Recurrence = scipy.sparse.linalg.cg as called by the reference (neural_cg/utils/validate.py:163-201):
//   r = b, x = 0; loop { if ||r|| < rtol*||b|| stop; z = M r; rho = r.z; p = z + (rho/rho_prev) p;
//                        q = A p; alpha = rho / p.q; x += alpha p; r -= alpha q }
// One iteration is 5 kernels for the SPAI preconditioner (4 for none/diagonal):
//   K1  t = G^T r                       (SpMV over the precomputed transposed layout, no atomics)
//   K2  z = G t + eps r ; rho = r.z     (SpMV, fused axpy + dot; last CTA: beta = rho/rho_prev)
//   K3  p = z + beta p
//   K4  q = A p ; pq = p.q              (SpMV, fused dot; last CTA: alpha = rho/pq)
//   K5  x += alpha p ; r -= alpha q ; rr = r.r   (last CTA: iter++, history, convergence flag)
// No scalar ever visits the host during the loop.  Two drivers:
//   mode 1 (stream): iterations are enqueued in chunks; every kernel starts with `if (done) return`,
//           and the host looks at a done flag copied to pinned memory two chunks behind the GPU.
//   mode 2 (graph):  one CUDA graph = init kernel + a WHILE conditional node whose body is the 5
//           kernels; K5's last CTA sets the loop condition (cudaGraphSetConditional).  One launch
//           per solve, no host involvement until the result is read.
#include <cstdlib>
#include <vector>

#include "csr.h"
#include <mutex>
#include <vector>

#include "pcg_kernels.cuh"
#include "pcg_small.cuh"
#include "pcg_small2.cuh"
#include "pcg_fused.cuh"

using namespace b200spai;

struct b200spai_pcg_s {
  b200spai_csr_s* A = nullptr;
  b200spai_csr_s* G = nullptr;
  b200spai_csr_s* GT = nullptr;  // owned
  double eps = 0.0;
  int n = 0;
  double *r = nullptr, *z = nullptr, *p = nullptr, *q = nullptr, *t = nullptr, *d = nullptr, *bw = nullptr, *xw = nullptr;
  bool have_d = false;
  PcgScalars* sc = nullptr;
  double* partials = nullptr;
  unsigned* ticket = nullptr;
  double* hist = nullptr;
  long long hist_cap = 0;
  PcgScalars* h_in = nullptr;   // pinned
  PcgScalars* h_out = nullptr;  // pinned
  int* h_done = nullptr;        // pinned [2]
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_chunk[2] = {nullptr, nullptr};
  cudaStream_t body_stream = nullptr, cap_stream = nullptr;  // graphs are captured on private streams
  int scaffold_dev = 0;
  int mode = 0;
  long long launches = 0;
  float setup_ms = 0.f;
  cudaGraphExec_t exec[4] = {nullptr, nullptr, nullptr, nullptr};
  cudaGraph_t graph[4] = {nullptr, nullptr, nullptr, nullptr};
  unsigned long long cond[4] = {0, 0, 0, 0};
  int launches_per_iter[4] = {0, 0, 0, 0};
  // fused persistent kernel (mode 4)
  FusedSync* fsync = nullptr;
  unsigned long long* fprof = nullptr;  // [4] device
  FusedPlan fplan;
  bool fplan_valid = false;
};

namespace b200spai {

// Host-side scaffolding of a solver (three small pinned buffers, four events, two capture streams): creating it costs
// ~1.7 ms, which is most of a small system's setup, so destroyed solvers park theirs for the next pcg_create.
struct PcgScaffold {
  int dev = 0;
  PcgScalars *h_in = nullptr, *h_out = nullptr;
  int* h_done = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_chunk[2] = {nullptr, nullptr};
  cudaStream_t body_stream = nullptr, cap_stream = nullptr;
};
static std::mutex g_scaffold_mu;
static std::vector<PcgScaffold>& scaffold_cache() {
  static auto* v = new std::vector<PcgScaffold>();  // leaked on purpose (see the memory pool)
  return *v;
}

static int scaffold_acquire(b200spai_pcg_s* s) {
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  PcgScaffold f;
  bool hit = false;
  {
    std::lock_guard<std::mutex> lock(g_scaffold_mu);
    auto& c = scaffold_cache();
    for (size_t i = 0; i < c.size(); ++i)
      if (c[i].dev == dev) {
        f = c[i];
        c.erase(c.begin() + (long)i);
        hit = true;
        break;
      }
  }
  if (!hit) {
    f.dev = dev;
    B2_CUDA(cudaMallocHost((void**)&f.h_in, sizeof(PcgScalars)));
    B2_CUDA(cudaMallocHost((void**)&f.h_out, sizeof(PcgScalars)));
    B2_CUDA(cudaMallocHost((void**)&f.h_done, 2 * sizeof(int)));
    B2_CUDA(cudaEventCreate(&f.ev0));
    B2_CUDA(cudaEventCreate(&f.ev1));
    B2_CUDA(cudaEventCreateWithFlags(&f.ev_chunk[0], cudaEventDisableTiming));
    B2_CUDA(cudaEventCreateWithFlags(&f.ev_chunk[1], cudaEventDisableTiming));
    B2_CUDA(cudaStreamCreateWithFlags(&f.body_stream, cudaStreamNonBlocking));
    B2_CUDA(cudaStreamCreateWithFlags(&f.cap_stream, cudaStreamNonBlocking));
  }
  s->scaffold_dev = f.dev;
  s->h_in = f.h_in;
  s->h_out = f.h_out;
  s->h_done = f.h_done;
  s->ev0 = f.ev0;
  s->ev1 = f.ev1;
  s->ev_chunk[0] = f.ev_chunk[0];
  s->ev_chunk[1] = f.ev_chunk[1];
  s->body_stream = f.body_stream;
  s->cap_stream = f.cap_stream;
  return 0;
}

static void scaffold_release(b200spai_pcg_s* s) {
  if (!s->h_in) return;
  PcgScaffold f;
  f.dev = s->scaffold_dev;
  f.h_in = s->h_in;
  f.h_out = s->h_out;
  f.h_done = s->h_done;
  f.ev0 = s->ev0;
  f.ev1 = s->ev1;
  f.ev_chunk[0] = s->ev_chunk[0];
  f.ev_chunk[1] = s->ev_chunk[1];
  f.body_stream = s->body_stream;
  f.cap_stream = s->cap_stream;
  s->h_in = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_scaffold_mu);
    if (scaffold_cache().size() < 16 && f.h_out && f.h_done && f.ev0 && f.ev1 && f.ev_chunk[0] && f.ev_chunk[1] &&
        f.body_stream && f.cap_stream) {
      scaffold_cache().push_back(f);
      return;
    }
  }
  if (f.h_in) cudaFreeHost(f.h_in);
  if (f.h_out) cudaFreeHost(f.h_out);
  if (f.h_done) cudaFreeHost(f.h_done);
  if (f.ev0) cudaEventDestroy(f.ev0);
  if (f.ev1) cudaEventDestroy(f.ev1);
  if (f.ev_chunk[0]) cudaEventDestroy(f.ev_chunk[0]);
  if (f.ev_chunk[1]) cudaEventDestroy(f.ev_chunk[1]);
  if (f.body_stream) cudaStreamDestroy(f.body_stream);
  if (f.cap_stream) cudaStreamDestroy(f.cap_stream);
}

// auto mode: the graph driver, unless B200SPAI_PCG_MODE=1 asks for stream-enqueued kernels (ncu cannot see the
// kernel nodes of a graph that holds a conditional node, so the profiling passes run in mode 1)
static int default_mode() {
  static int v = 0;
  if (v == 0) {
    const char* e = getenv("B200SPAI_PCG_MODE");
    v = (e && atoi(e) == 1) ? 1 : 2;
  }
  return v;
}

static int ensure_hist(b200spai_pcg_s* s, long long need) {
  if (need <= s->hist_cap) return 0;
  dev_free(s->hist);
  s->hist = nullptr;
  B2_TRY(dev_alloc(&s->hist, (size_t)need));
  s->hist_cap = need;
  for (int i = 0; i < 4; ++i) {  // graphs bake the history pointer
    if (s->exec[i]) cudaGraphExecDestroy(s->exec[i]);
    if (s->graph[i]) cudaGraphDestroy(s->graph[i]);
    s->exec[i] = nullptr;
    s->graph[i] = nullptr;
  }
  return 0;
}

static int ensure_diag(b200spai_pcg_s* s, cudaStream_t st) {
  if (s->have_d) return 0;
  // the buffer is allocated once: captured graphs and the fused plan bake its address, a value refresh only rewrites it
  if (!s->d) B2_TRY(dev_alloc(&s->d, (size_t)s->n));
  B2_TRY(b200spai_csr_diagonal(s->A, s->d, st));
  s->have_d = true;
  return 0;
}

// z = M r (+ rho = r.z when sc != nullptr)
static int enqueue_precond(b200spai_pcg_s* s, int precond, const double* r, double* z, PcgScalars* sc, cudaStream_t st,
                           int* nlaunch) {
  if (precond == B200SPAI_PRECOND_NONE || precond == B200SPAI_PRECOND_DIAGONAL) {
    B2_ARG(sc != nullptr, "jacobi preconditioner needs the scalar block");
    pcg_jacobi_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(r, precond == B200SPAI_PRECOND_DIAGONAL ? s->d : nullptr, z,
                                                           s->n, sc, s->partials, s->ticket);
    B2_LAUNCH_CHECK();
    *nlaunch += 1;
    return 0;
  }
  const bool scaled = (precond == B200SPAI_PRECOND_EXT_SPAI_SCALED);
  SpmvArgs a1{};
  a1.x = r;
  a1.y = s->t;
  a1.d = scaled ? s->d : nullptr;
  a1.sc = sc;
  B2_TRY(spmv_launch(s->GT, a1, scaled ? SPMV_SCALE : SPMV_PLAIN, 0, st));
  SpmvArgs a2{};
  a2.x = s->t;
  a2.y = z;
  a2.aux = r;
  a2.d = scaled ? s->d : nullptr;
  a2.eps = s->eps;
  a2.partials = s->partials;
  a2.ticket = s->ticket;
  a2.sc = sc;
  a2.final_op = FINAL_RHO;
  B2_TRY(spmv_launch(s->G, a2, SPMV_AXPY_DOT, 0, st));
  *nlaunch += 2;
  return 0;
}

static int enqueue_iteration(b200spai_pcg_s* s, int precond, double* x, cudaStream_t st, int* nlaunch) {
  B2_TRY(enqueue_precond(s, precond, s->r, s->z, s->sc, st, nlaunch));
  pcg_update_p_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(s->z, s->p, s->n, s->sc);
  SpmvArgs a{};
  a.x = s->p;
  a.aux = s->p;
  a.y = s->q;
  a.partials = s->partials;
  a.ticket = s->ticket;
  a.sc = s->sc;
  a.final_op = FINAL_PQ;
  B2_TRY(spmv_launch(s->A, a, SPMV_DOT, 0, st));
  pcg_update_xr_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(x, s->r, s->p, s->q, s->n, s->sc, s->partials, s->ticket,
                                                            s->hist);
  B2_LAUNCH_CHECK();
  *nlaunch += 3;
  return 0;
}

static int build_graph(b200spai_pcg_s* s, int precond, cudaStream_t st) {
  if (s->exec[precond]) return 0;
  // make sure the lazily built partitions / compressed streams of EVERY matrix that takes the streaming kernel exist
  // before capture (their build launches kernels and synchronises, which would invalidate the capture)
  {
    SpmvArgs warm{};
    bool any = false;
    if (spmv_auto_kernel(s->A) == 2) {
      warm.x = s->p;
      warm.y = s->q;
      B2_TRY(spmv_launch(s->A, warm, SPMV_PLAIN, 0, st));
      any = true;
    }
    if (precond >= B200SPAI_PRECOND_EXT_SPAI) {
      if (spmv_auto_kernel(s->GT) == 2) {
        warm.x = s->r;
        warm.y = s->t;
        B2_TRY(spmv_launch(s->GT, warm, SPMV_PLAIN, 0, st));
        any = true;
      }
      if (spmv_auto_kernel(s->G) == 2) {
        warm.x = s->t;
        warm.y = s->z;
        B2_TRY(spmv_launch(s->G, warm, SPMV_PLAIN, 0, st));
        any = true;
      }
    }
    if (any) B2_CUDA(cudaStreamSynchronize(st));
  }
  cudaGraph_t graph = nullptr;
  cudaStream_t cs = s->cap_stream;  // the caller's stream may be the legacy default stream (not capturable)
  B2_CUDA(cudaStreamBeginCapture(cs, cudaStreamCaptureModeRelaxed));
  pcg_init_kernel<<<vec_grid(s->n), VTHREADS, 0, cs>>>(s->bw, s->xw, s->r, s->n, s->sc, s->partials, s->ticket, s->hist);
  cudaStreamCaptureStatus status;
  const cudaGraphNode_t* deps = nullptr;
  size_t ndeps = 0;
  B2_CUDA(cudaStreamGetCaptureInfo(cs, &status, nullptr, &graph, &deps, &ndeps));
  cudaGraphConditionalHandle handle;
  B2_CUDA(cudaGraphConditionalHandleCreate(&handle, graph, 1, cudaGraphCondAssignDefault));
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
  int rc = enqueue_iteration(s, precond, s->xw, s->body_stream, &nl);
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
  s->cond[precond] = (unsigned long long)handle;
  s->launches_per_iter[precond] = nl;
  return 0;
}

// Small systems: the whole solve in one cooperative kernel (pcg_small.cuh).  Eligible when every row can be given a
// thread of a co-resident grid a few times over and the device supports cooperative launches.
static int small_grid(const b200spai_pcg_s* s) {
  static int occ = -1, coop = -1;
  if (occ < 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pcg_small_kernel, SMALL_THREADS, 0) != cudaSuccess) occ = 0;
  }
  if (!coop || occ < 1 || s->n < 1 || s->n > 131072) return 0;
  long long g = ((long long)s->n + SMALL_THREADS - 1) / SMALL_THREADS;
  const long long cap = std::min<long long>((long long)sm_count() * occ, SMALL_MAX_CTAS);
  if (g > cap) g = cap;
  return (int)g;
}

// Second-generation small-system kernel (pcg_small2.cuh): L lanes per row, LL all-reduce.  B200SPAI_SMALL_KERNEL=1 keeps
// the thread-per-row kernel (development switch for A/B measurements).
static bool small_v1() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200SPAI_SMALL_KERNEL");
    v = (e && atoi(e) == 1) ? 1 : 0;
  }
  return v == 1;
}

// Longest row of a matrix (cached in the handle): decides between the two small-system kernels.
static __global__ void max_row_len_kernel(const int* __restrict__ rowptr, int n, int* __restrict__ out) {
  int m = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) m = max(m, rowptr[i + 1] - rowptr[i]);
  m = __reduce_max_sync(0xffffffffu, m);
  if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}
static int csr_max_row_len(b200spai_csr_s* A, cudaStream_t st, int* out) {
  if (A->max_row_len < 0) {
    int* d = nullptr;
    B2_TRY(dev_alloc(&d, 1));
    B2_CUDA(cudaMemsetAsync(d, 0, sizeof(int), st));
    if (A->n_rows > 0) max_row_len_kernel<<<std::min(64, (A->n_rows + 255) / 256), 256, 0, st>>>(A->rowptr, A->n_rows, d);
    int h = 0;
    B2_CUDA(cudaMemcpyAsync(&h, d, sizeof(int), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    dev_free(d);
    A->max_row_len = h;
    count_launches(1);
  }
  *out = A->max_row_len;
  return 0;
}

// Plan of a small2 solve: lanes per row, grid, dynamic shared memory.  ok = false: the matrices do not fit into the shared
// memory of a co-resident grid (e.g. the 49 152-row elasticity system: 67 MB) -- the thread-per-row kernel takes it.
struct Small2Plan {
  bool ok = false;
  int lg = 0, grid = 0, smem = 0;
};

template <int LG>
static int small2_occ(int smem, int* occ) {
  // the opt-in limit is a per-function, process-wide setting: raised once to the maximum (concurrent solves on other
  // threads launch with other sizes), the size of a launch is passed with the launch
  static std::once_flag once;
  static cudaError_t err = cudaSuccess;
  std::call_once(once, [] { err = cudaFuncSetAttribute(pcg_small2_kernel<LG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 1024); });
  B2_CUDA(err);
  B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, pcg_small2_kernel<LG>, S2_THREADS, smem));
  return 0;
}
static int small2_occ_lg(int lg, int smem, int* occ) {
  switch (lg) {
    case 0: return small2_occ<0>(smem, occ);
    case 1: return small2_occ<1>(smem, occ);
    case 2: return small2_occ<2>(smem, occ);
    case 3: return small2_occ<3>(smem, occ);
    case 4: return small2_occ<4>(smem, occ);
    default: return small2_occ<5>(smem, occ);
  }
}
template <int LG>
static int small2_launch(const Small2Args& a, const Small2Plan& pl, cudaStream_t st) {
  B2_CUDA(cudaMemsetAsync(a.slots, 0, (size_t)2 * pl.grid * 4 * sizeof(unsigned long long), st));  // no tag of an earlier solve
  void* args[] = {(void*)&a};
  B2_CUDA(cudaLaunchCooperativeKernel((const void*)pcg_small2_kernel<LG>, dim3((unsigned)pl.grid), dim3(S2_THREADS), args,
                                      (size_t)pl.smem, st));
  return 0;
}

static int small2_plan(const b200spai_pcg_s* s, int precond, Small2Plan* pl) {
  *pl = Small2Plan{};
  const int n = s->n;
  if (n < 1 || n > 131072) return 0;
  const bool spai = precond >= B200SPAI_PRECOND_EXT_SPAI;
  long long nnz_max = s->A->nnz, nnz_all = s->A->nnz;
  if (spai) {
    nnz_max = std::max(nnz_max, s->G->nnz);
    nnz_all += s->G->nnz + s->GT->nnz;
  }
  const int nmat = spai ? 3 : 1;
  // lanes per row: as many as one round of 8 passes needs for a typical row, as few as keep (about) one row per group
  const double mean = (double)nnz_max / (double)n;
  int lg = 0;
  while (lg < 5 && 8.0 * (1 << lg) < mean) ++lg;  // (this kernel takes single systems with long / irregular rows: measured, one lane per row loses there)
  const long long lanes_cap = 2LL * sm_count() * S2_THREADS;
  while (lg > 0 && (long long)n * (1 << lg) > lanes_cap) --lg;
  const long long ng = S2_THREADS >> lg;
  const long long g_max = std::min<long long>(2LL * sm_count(), S2_MAX_CTAS);
  const long long g_rows = std::min<long long>(((long long)n + ng - 1) / ng, g_max);
  // shared memory per CTA: its share of the matrices (12 bytes per entry + row pointers), 12 % head room for uneven rows
  // (a CTA whose rows of a matrix do not fit reads that matrix from global memory).  More CTAs than the rows need are
  // taken when that is what makes the share fit (some row groups then idle).
  const int per_sm = 227 * 1024 - 2048;
  long long g = 0, want = 0;
  for (long long cand : {g_rows, g_max, std::max<long long>(g_rows, (long long)sm_count())}) {
    const double share = ((double)nnz_all * 12.0 + (double)nmat * 4.0 * (double)(n + cand)) / (double)cand;
    const long long w = (long long)(share * 1.12) + 48LL * nmat + 1024;
    const long long cap = (cand > sm_count()) ? per_sm / 2 - 1024 : per_sm;
    if (w <= cap) {
      g = cand;
      want = w;
      break;
    }
  }
  if (g == 0) return 0;
  int smem = (int)((want + 127) & ~127LL);
  int occ = 0;
  B2_TRY(small2_occ_lg(lg, smem, &occ));
  if ((long long)occ * sm_count() < g) return 0;
  pl->ok = true;
  pl->lg = lg;
  pl->grid = (int)g;
  pl->smem = smem;
  return 0;
}

static int solve_small2(b200spai_pcg_s* s, int precond, const double* b, double* x, const Small2Plan& pl, cudaStream_t st) {
  Small2Args a{};
  a.A = s->A->view();
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) {
    a.G = s->G->view();
    a.GT = s->GT->view();
  }
  a.b = b;
  a.x = x;
  a.r0 = s->r;
  a.r1 = s->bw;
  a.p0 = s->p;
  a.p1 = s->xw;
  a.z = s->z;
  a.q = s->q;
  a.t = s->t;
  a.d = (precond == B200SPAI_PRECOND_DIAGONAL || precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) ? s->d : nullptr;
  a.eps = s->eps;
  a.precond = precond;
  a.sc = s->sc;
  a.slots = reinterpret_cast<unsigned long long*>(s->partials);
  a.hist = s->hist;
  a.smem_bytes = pl.smem;
  static_assert(2 * S2_MAX_CTAS * 4 <= MAX_PARTIALS, "two generations of LL slots must fit the partials buffer");
  int khz = 1965000, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  a.spin_limit = (long long)(20.0 * 1e3 * (double)khz);
  switch (pl.lg) {
    case 0: B2_TRY(small2_launch<0>(a, pl, st)); break;
    case 1: B2_TRY(small2_launch<1>(a, pl, st)); break;
    case 2: B2_TRY(small2_launch<2>(a, pl, st)); break;
    case 3: B2_TRY(small2_launch<3>(a, pl, st)); break;
    case 4: B2_TRY(small2_launch<4>(a, pl, st)); break;
    default: B2_TRY(small2_launch<5>(a, pl, st)); break;
  }
  s->launches += 1;
  count_launches(1);
  return 0;
}

static int solve_small(b200spai_pcg_s* s, int precond, const double* b, double* x, int grid, cudaStream_t st) {
  if (!small_v1()) {
    // Measured (profiles/r02_shapes_small_kernels.txt): the lanes-per-row kernel wins where rows are long or of very
    // different length (synthetic SPD: -25 % per iteration); on mesh matrices (15 +- 2 entries per row) the
    // thread-per-row kernel is as fast (heat: 19.8 vs 21.6 us per iteration) and leaves the shared memory to other solves
    static const int force = [] { const char* e = getenv("B200SPAI_SMALL_KERNEL"); return (e && atoi(e) == 2) ? 1 : 0; }();
    b200spai_csr_s* W = (precond >= B200SPAI_PRECOND_EXT_SPAI && s->G->nnz > s->A->nnz) ? s->G : s->A;
    int longest = 0;
    B2_TRY(csr_max_row_len(W, st, &longest));
    const double mean = (double)W->nnz / (double)std::max(1, s->n);
    if (force || mean > 16.0 || (double)longest >= 2.5 * mean) {
      Small2Plan pl;
      B2_TRY(small2_plan(s, precond, &pl));
      if (pl.ok) return solve_small2(s, precond, b, x, pl, st);
    }
  }
  SmallArgs a{};
  a.A = s->A->view();
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) {
    a.G = s->G->view();
    a.GT = s->GT->view();
  }
  a.b = b;
  a.x = x;
  a.r0 = s->r;
  a.r1 = s->bw;
  a.p0 = s->p;
  a.p1 = s->xw;
  a.z = s->z;
  a.q = s->q;
  a.t = s->t;
  a.d = (precond == B200SPAI_PRECOND_DIAGONAL || precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) ? s->d : nullptr;
  a.eps = s->eps;
  a.precond = precond;
  a.sc = s->sc;
  a.partials = s->partials;
  a.hist = s->hist;
  static_assert(6 * SMALL_MAX_CTAS <= MAX_PARTIALS, "three two-value reduction slots must fit the partials buffer");
  void* args[] = {&a};
  B2_CUDA(cudaLaunchCooperativeKernel((const void*)pcg_small_kernel, dim3((unsigned)grid), dim3(SMALL_THREADS), args, 0, st));
  s->launches += 1;
  count_launches(1);
  return 0;
}

// Large systems with a SPAI preconditioner: the whole solve in the persistent fused kernel (pcg_fused.cuh).
static int fused_plan_for(b200spai_pcg_s* s, cudaStream_t st) {
  if (!s->fplan_valid) {
    B2_TRY(fused_plan(s->A, s->G, s->GT, st, &s->fplan));
    s->fplan_valid = true;
  }
  return 0;
}

static int solve_fused(b200spai_pcg_s* s, int precond, const double* b, double* x, cudaStream_t st) {
  if (!s->fsync) B2_TRY(dev_alloc(&s->fsync, 1));
  if (!s->fprof) B2_TRY(dev_alloc(&s->fprof, 10));
  B2_CUDA(cudaMemsetAsync(s->fsync, 0, sizeof(FusedSync), st));
  B2_CUDA(cudaMemsetAsync(s->fprof, 0, 10 * sizeof(unsigned long long), st));
  FusedArgs a{};
  a.A = fused_mat(s->A, s->fplan);
  a.G = fused_mat(s->G, s->fplan);
  a.GT = fused_mat(s->GT, s->fplan);
  a.n = s->n;
  a.off = 0;
  a.hi_halo = 0;
  a.b = b;
  a.x = x;
  a.r = s->r;
  a.p = s->p;
  a.z = s->z;
  a.q = s->q;
  a.t = s->t;
  a.d = (precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) ? s->d : nullptr;
  a.eps = s->eps;
  a.precond = precond;
  a.sc = s->sc;
  a.partials = s->partials;
  a.hist = s->hist;
  a.sync = s->fsync;
  a.prof = s->fprof;
  int khz = 1965000, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  a.spin_limit = (long long)(20.0 * 1e3 * (double)khz);  // 20 s: a CTA that never arrives ends the solve instead of hanging the GPU
  a.peer.rank = 0;
  a.peer.world = 1;
  a.peer.n_push = 0;
  a.peer.n_recv = 0;
  B2_TRY(fused_launch(a, s->fplan, st));
  s->launches += 1;
  count_launches(1);
  return 0;
}

static int solve_impl(b200spai_pcg_s* s, int precond, const double* b, double* x, double rtol, int max_iter,
                      double* hist_out, cudaStream_t st, b200spai_pcg_result* res) {
  B2_ARG(s && b && x && res, "pcg_solve: NULL argument");
  B2_ARG(precond >= 0 && precond <= 3, "pcg_solve: unknown preconditioner");
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) B2_ARG(s->G && s->GT, "pcg_solve: ext_spai needs b200spai_pcg_set_spai first");
  if (precond == B200SPAI_PRECOND_DIAGONAL || precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) B2_TRY(ensure_diag(s, st));
  if (max_iter <= 0) max_iter = s->n;
  B2_TRY(ensure_hist(s, (long long)max_iter + 1));
  int mode = s->mode;
  const int sgrid = (mode == 0 || mode == 3) ? small_grid(s) : 0;
  if (mode == 3) B2_ARG(sgrid > 0, "pcg_solve: mode 3 (single cooperative kernel) needs a system of <= 131072 rows");
  if (mode == 0) mode = (sgrid > 0 && !getenv("B200SPAI_NO_SMALL") && default_mode() == 2) ? 3 : default_mode();
  // auto takes the fused persistent kernel for large SPAI solves only when B200SPAI_PCG_MODE=4 asks for it: on ONE GPU
  // it runs the same sweeps as the graph driver, whose launch gaps are already < 1 % of an iteration at these sizes
  static const bool auto_fused = [] { const char* e = getenv("B200SPAI_PCG_MODE"); return e && atoi(e) == 4; }();
  if ((mode == 2 && s->mode == 0 && auto_fused) || mode == 4) {
    bool fused = false;
    if (precond >= B200SPAI_PRECOND_EXT_SPAI) {
      B2_TRY(fused_plan_for(s, st));
      fused = s->fplan.which >= 0;
    }
    if (mode == 4) B2_ARG(fused, "pcg_solve: mode 4 (fused persistent kernel) needs a SPAI preconditioner and matrices that share one stream format");
    if (fused) mode = 4;
  }

  PcgScalars init{};
  init.rtol = rtol;
  init.max_iter = max_iter;
  if (mode == 2) {
    B2_TRY(build_graph(s, precond, st));
    init.use_cond = 1;
    init.cond_handle = s->cond[precond];
  }
  *s->h_in = init;
  B2_CUDA(cudaEventRecord(s->ev0, st));
  B2_CUDA(cudaMemcpyAsync(s->sc, s->h_in, sizeof(PcgScalars), cudaMemcpyHostToDevice, st));
  if (mode == 3) {
    B2_TRY(solve_small(s, precond, b, x, sgrid, st));
  } else if (mode == 4) {
    B2_TRY(solve_fused(s, precond, b, x, st));
  } else if (mode == 2) {
    B2_CUDA(cudaMemcpyAsync(s->bw, b, (size_t)s->n * sizeof(double), cudaMemcpyDeviceToDevice, st));
    B2_CUDA(cudaGraphLaunch(s->exec[precond], st));
    B2_CUDA(cudaMemcpyAsync(x, s->xw, (size_t)s->n * sizeof(double), cudaMemcpyDeviceToDevice, st));
  } else {
    double* xs = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) ? x : s->xw;  // the vector kernels use 16-byte accesses
    pcg_init_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(b, xs, s->r, s->n, s->sc, s->partials, s->ticket, s->hist);
    B2_LAUNCH_CHECK();
    { const long long _nl = (1); s->launches += _nl; count_launches(_nl); }
    const int CH = 8;
    int slot = 0;
    long long enq = 0;
    bool first = true;
    s->h_done[0] = s->h_done[1] = 0;
    while (true) {
      int nl = 0;
      for (int c = 0; c < CH; ++c) B2_TRY(enqueue_iteration(s, precond, xs, st, &nl));
      { const long long _nl = (nl); s->launches += _nl; count_launches(_nl); }
      enq += CH;
      B2_CUDA(cudaMemcpyAsync(s->h_done + slot, &s->sc->done, sizeof(int), cudaMemcpyDeviceToHost, st));
      B2_CUDA(cudaEventRecord(s->ev_chunk[slot], st));
      slot ^= 1;
      if (first) {
        first = false;
        if (enq < (long long)max_iter + CH) continue;
      }
      B2_CUDA(cudaEventSynchronize(s->ev_chunk[slot]));
      if (s->h_done[slot]) break;
      if (enq >= (long long)max_iter + 2 * CH) break;
    }
    if (xs != x) B2_CUDA(cudaMemcpyAsync(x, xs, (size_t)s->n * sizeof(double), cudaMemcpyDeviceToDevice, st));
  }
  B2_CUDA(cudaEventRecord(s->ev1, st));
  B2_CUDA(cudaMemcpyAsync(s->h_out, s->sc, sizeof(PcgScalars), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  const PcgScalars& o = *s->h_out;
  if (mode == 2) { const long long _nl = (1 + (long long)s->launches_per_iter[precond] * (o.iter > 0 ? o.iter : 1)); s->launches += _nl; count_launches(_nl); }
  res->iterations = o.iter;
  res->converged = o.converged;
  res->breakdown = o.breakdown;
  res->reserved = mode;
  res->b_norm = o.bnorm;
  res->final_residual = sqrt(o.rr);
  res->setup_ms = s->setup_ms;
  B2_CUDA(cudaEventElapsedTime(&res->solve_ms, s->ev0, s->ev1));
  if (hist_out)
    B2_CUDA(cudaMemcpyAsync(hist_out, s->hist, ((size_t)o.iter + 1) * sizeof(double), cudaMemcpyDefault, st));
  B2_CUDA(cudaStreamSynchronize(st));
  return 0;
}

}  // namespace b200spai

extern "C" {

int b200spai_pcg_create(b200spai_csr_t a, b200spai_pcg_t* out) {
  B2_ARG(a && out, "pcg_create: NULL argument");
  B2_ARG(a->n_rows == a->n_cols, "pcg_create: A must be square");
  auto* s = new b200spai_pcg_s();
  s->A = a;
  s->n = a->n_rows;
  const size_t n = (size_t)s->n;
  int rc = 0;
  for (double** v : {&s->r, &s->z, &s->p, &s->q, &s->t, &s->bw, &s->xw})
    if (!rc) rc = dev_alloc(v, n);
  if (!rc) rc = dev_alloc(&s->sc, 1);
  if (!rc) rc = dev_alloc(&s->partials, MAX_PARTIALS);
  if (!rc) rc = dev_alloc(&s->ticket, 1);
  if (rc) {
    b200spai_pcg_destroy(s);
    return rc;
  }
  B2_CUDA(cudaMemset(s->ticket, 0, sizeof(unsigned)));
  B2_CUDA(cudaMemset(s->sc, 0, sizeof(PcgScalars)));
  B2_CUDA(cudaMemset(s->p, 0, n * sizeof(double)));
  // legacy-default-stream memsets do not order with a caller's non-blocking stream: finish them here, or the zeroing
  // of p could land in the middle of the first solve
  B2_CUDA(cudaStreamSynchronize(cudaStreamLegacy));
  if (scaffold_acquire(s)) {
    b200spai_pcg_destroy(s);
    return B200SPAI_ERR_CUDA;
  }
  *out = s;
  return 0;
}

int b200spai_pcg_set_spai(b200spai_pcg_t s, b200spai_csr_t g, double epsilon, void* stream) {
  B2_ARG(s && g, "pcg_set_spai: NULL argument");
  B2_ARG(g->n_rows == s->n && g->n_cols == s->n, "pcg_set_spai: G must have A's shape");
  cudaStream_t st = as_stream(stream);
  cudaEvent_t e0, e1;
  B2_CUDA(cudaEventCreate(&e0));
  B2_CUDA(cudaEventCreate(&e1));
  B2_CUDA(cudaEventRecord(e0, st));
  if (s->GT) csr_free(s->GT);
  s->GT = nullptr;
  for (int i = 2; i < 4; ++i) {  // SPAI graphs bake G's pointers
    if (s->exec[i]) cudaGraphExecDestroy(s->exec[i]);
    if (s->graph[i]) cudaGraphDestroy(s->graph[i]);
    s->exec[i] = nullptr;
    s->graph[i] = nullptr;
  }
  // G built with a refresh plan (fixed topology): G^T keeps its permutation too, so that b200spai_pcg_refresh can
  // re-gather its values instead of sorting 150 M keys again
  if (g->src_kind == 1) B2_TRY(b200spai_csr_transpose_plan(g, stream, &s->GT));
  else B2_TRY(b200spai_csr_transpose(g, stream, &s->GT));
  s->fplan_valid = false;
  s->G = g;
  s->eps = epsilon;
  B2_CUDA(cudaEventRecord(e1, st));
  B2_CUDA(cudaEventSynchronize(e1));
  B2_CUDA(cudaEventElapsedTime(&s->setup_ms, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return 0;
}

int b200spai_pcg_refresh(b200spai_pcg_t s, void* stream) {
  B2_ARG(s, "pcg_refresh: NULL handle");
  if (s->G) {
    B2_ARG(s->GT && (s->GT->nnz == 0 || s->GT->src_kind == 2),
           "pcg_refresh: G^T has no refresh plan (G must come from b200spai_edges_to_csr_plan before b200spai_pcg_set_spai)");
    B2_TRY(b200spai_csr_refresh_transposed(s->GT, s->G, stream));
  }
  s->have_d = false;  // diag(A) is re-read on the next diagonal / scaled solve
  return 0;
}

int b200spai_pcg_set_mode(b200spai_pcg_t s, int mode) {
  B2_ARG(s && mode >= 0 && mode <= 4, "pcg_set_mode: mode must be 0, 1, 2, 3 or 4");
  s->mode = mode;
  return 0;
}

int b200spai_pcg_solve_dev(b200spai_pcg_t s, int precond, const double* b, double* x, double rtol, int32_t max_iter,
                           double* hist, void* stream, b200spai_pcg_result* result) {
  return solve_impl(s, precond, b, x, rtol, max_iter, hist, as_stream(stream), result);
}

// Many independent small systems in waves of ONE launch each (pcg_small2.cuh, S2Sys descriptors): the solver is bound to
// the block-diagonal batch matrix (a PyG-style batch: the reference collates samples this way and splits them by `ptr`
// before solving, workspace.py:137-143); block s = rows [ptr[s], ptr[s+1]) is solved by its own CTA group with its own
// scalars, stopping rule and iteration count -- the arithmetic of a stand-alone solve of that block.
int b200spai_pcg_solve_batch(b200spai_pcg_t s, int precond, int32_t n_sys, const int32_t* row_ptr_host, const double* b,
                             double* x, double rtol, int32_t max_iter, double* hist, int64_t hist_stride, void* stream,
                             b200spai_pcg_result* results) {
  B2_ARG(s && row_ptr_host && b && x && results && n_sys >= 1, "pcg_solve_batch: bad argument");
  B2_ARG(precond >= 0 && precond <= 3, "pcg_solve_batch: unknown preconditioner");
  B2_ARG(row_ptr_host[0] == 0 && row_ptr_host[n_sys] == s->n, "pcg_solve_batch: row_ptr must run from 0 to the number of rows");
  B2_ARG(!hist || hist_stride > 0, "pcg_solve_batch: a history buffer needs its per-system stride");
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) B2_ARG(s->G && s->GT, "pcg_solve_batch: ext_spai needs b200spai_pcg_set_spai first");
  cudaStream_t st = as_stream(stream);
  if (precond == B200SPAI_PRECOND_DIAGONAL || precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) B2_TRY(ensure_diag(s, st));
  const bool spai = precond >= B200SPAI_PRECOND_EXT_SPAI;
  long long nnz_max = s->A->nnz, nnz_all = s->A->nnz;
  if (spai) {
    nnz_max = std::max(nnz_max, s->G->nnz);
    nnz_all += s->G->nnz + s->GT->nnz;
  }
  const double mean = (double)nnz_max / (double)std::max(1, s->n);
  int lg = 0;
  while (lg < 5 && (lg == 0 ? 16.0 : 8.0 * (1 << lg)) < mean) ++lg;  // entries per round: 16 with one lane per row, else 8 L
  int occ = 0;
  B2_TRY(small2_occ_lg(lg, 0, &occ));
  const int cap = std::min<int>(sm_count() * std::min(occ, 2), S2_MAX_CTAS);
  B2_ARG(cap >= 1, "pcg_solve_batch: the kernel does not fit this device");
  // waves: systems whose matrices stay L2-resident together (they are re-read every iteration), at least 2 CTAs each
  const double bytes_per_row = 12.0 * (double)nnz_all / (double)std::max(1, s->n) + 64.0;
  const double wave_bytes = 88e6;
  std::vector<PcgScalars> h_sc((size_t)n_sys);
  std::vector<S2Sys> h_sys((size_t)n_sys);
  PcgScalars* d_sc = nullptr;
  S2Sys* d_sys = nullptr;
  B2_TRY(dev_alloc(&d_sc, (size_t)n_sys));
  int rc = dev_alloc(&d_sys, (size_t)n_sys);
  if (rc) {
    dev_free(d_sc);
    return rc;
  }
  for (int i = 0; i < n_sys; ++i) {
    const int rows = row_ptr_host[i + 1] - row_ptr_host[i];
    if (rows < 0) {
      dev_free(d_sc);
      dev_free(d_sys);
      set_error("pcg_solve_batch: row_ptr must be non-decreasing");
      return B200SPAI_ERR_ARG;
    }
    PcgScalars init{};
    init.rtol = rtol;
    init.max_iter = max_iter > 0 ? max_iter : std::max(rows, 1);
    h_sc[(size_t)i] = init;
  }
  auto fail = [&](int code) {
    dev_free(d_sc);
    dev_free(d_sys);
    return code;
  };
  if (cudaMemcpyAsync(d_sc, h_sc.data(), sizeof(PcgScalars) * (size_t)n_sys, cudaMemcpyHostToDevice, st) != cudaSuccess)
    return fail(B200SPAI_ERR_CUDA);
  int khz = 1965000, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  B2_CUDA(cudaEventRecord(s->ev0, st));
  int first = 0, waves = 0;
  while (first < n_sys) {
    int last = first;
    double bytes = 0.0;
    while (last < n_sys && (last - first) < cap / 2 + (cap == 1) &&
           (last == first || bytes + bytes_per_row * (row_ptr_host[last + 1] - row_ptr_host[last]) <= wave_bytes)) {
      bytes += bytes_per_row * (row_ptr_host[last + 1] - row_ptr_host[last]);
      ++last;
    }
    const int cnt = last - first;
    const long long rows_wave = std::max(1, row_ptr_host[last] - row_ptr_host[first]);
    // CTAs in proportion to the rows, at least one each, never more than one row group per CTA would use
    const long long ng = S2_THREADS >> lg;
    int used = 0;
    for (int i = first; i < last; ++i) {
      const int rows = row_ptr_host[i + 1] - row_ptr_host[i];
      long long c = (long long)((double)cap * (double)rows / (double)rows_wave);
      c = std::max<long long>(1, std::min<long long>(c, (rows + ng - 1) / ng));
      h_sys[(size_t)i] = S2Sys{row_ptr_host[i], rows, 0, (int)c};
      used += (int)c;
    }
    while (used > cap) {  // rounding: take CTAs back from the largest groups
      int big = first;
      for (int i = first; i < last; ++i)
        if (h_sys[(size_t)i].ncta > h_sys[(size_t)big].ncta) big = i;
      if (h_sys[(size_t)big].ncta <= 1) break;
      h_sys[(size_t)big].ncta -= 1;
      used -= 1;
    }
    if (used > cap) return fail((set_error("pcg_solve_batch: more systems in a wave than co-resident CTAs"), B200SPAI_ERR_ARG));
    int c0 = 0;
    for (int i = first; i < last; ++i) {
      h_sys[(size_t)i].cta0 = c0;
      c0 += h_sys[(size_t)i].ncta;
    }
    if (cudaMemcpyAsync(d_sys + first, h_sys.data() + first, sizeof(S2Sys) * (size_t)cnt, cudaMemcpyHostToDevice, st) != cudaSuccess)
      return fail(B200SPAI_ERR_CUDA);
    Small2Args a{};
    a.A = s->A->view();
    if (spai) {
      a.G = s->G->view();
      a.GT = s->GT->view();
    }
    a.b = b;
    a.x = x;
    a.r0 = s->r;
    a.r1 = s->bw;
    a.p0 = s->p;
    a.p1 = s->xw;
    a.z = s->z;
    a.q = s->q;
    a.t = s->t;
    a.d = (precond == B200SPAI_PRECOND_DIAGONAL || precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) ? s->d : nullptr;
    a.eps = s->eps;
    a.precond = precond;
    a.sc = d_sc + first;
    a.slots = reinterpret_cast<unsigned long long*>(s->partials);
    a.hist = hist ? hist + (long long)first * hist_stride : nullptr;
    a.hist_stride = hist_stride;
    a.smem_bytes = 0;  // the matrices of a wave (tens of MB) live in L2, not in shared memory
    a.sys = d_sys + first;
    a.n_sys = cnt;
    a.spin_limit = (long long)(20.0 * 1e3 * (double)khz);
    Small2Plan pl;
    pl.ok = true;
    pl.lg = lg;
    pl.grid = used;
    pl.smem = 0;
    int lrc = 0;
    switch (lg) {
      case 0: lrc = small2_launch<0>(a, pl, st); break;
      case 1: lrc = small2_launch<1>(a, pl, st); break;
      case 2: lrc = small2_launch<2>(a, pl, st); break;
      case 3: lrc = small2_launch<3>(a, pl, st); break;
      case 4: lrc = small2_launch<4>(a, pl, st); break;
      default: lrc = small2_launch<5>(a, pl, st); break;
    }
    if (lrc) return fail(lrc);
    s->launches += 1;
    count_launches(1);
    first = last;
    ++waves;
  }
  B2_CUDA(cudaEventRecord(s->ev1, st));
  if (cudaMemcpyAsync(h_sc.data(), d_sc, sizeof(PcgScalars) * (size_t)n_sys, cudaMemcpyDeviceToHost, st) != cudaSuccess)
    return fail(B200SPAI_ERR_CUDA);
  if (cudaStreamSynchronize(st) != cudaSuccess) {
    set_error(std::string("pcg_solve_batch: ") + cudaGetErrorString(cudaGetLastError()));
    return fail(B200SPAI_ERR_CUDA);
  }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, s->ev0, s->ev1);
  for (int i = 0; i < n_sys; ++i) {
    const PcgScalars& o = h_sc[(size_t)i];
    b200spai_pcg_result& r = results[i];
    r.iterations = o.iter;
    r.converged = o.converged;
    r.breakdown = o.breakdown;
    r.reserved = 3;
    r.b_norm = o.bnorm;
    r.final_residual = sqrt(o.rr);
    r.setup_ms = s->setup_ms;
    r.solve_ms = ms;  // the whole batch (the systems of a wave run concurrently)
  }
  dev_free(d_sc);
  dev_free(d_sys);
  return 0;
}

int b200spai_pcg_solve_host(b200spai_pcg_t s, int precond, const double* b_host, double* x_host, double rtol,
                            int32_t max_iter, double* hist_host, void* stream, b200spai_pcg_result* result) {
  B2_ARG(s && b_host && x_host && result, "pcg_solve_host: NULL argument");
  cudaStream_t st = as_stream(stream);
  double *bd = nullptr, *xd = nullptr;
  B2_TRY(dev_alloc(&bd, (size_t)s->n));
  B2_TRY(dev_alloc(&xd, (size_t)s->n));
  B2_CUDA(cudaMemcpyAsync(bd, b_host, (size_t)s->n * sizeof(double), cudaMemcpyHostToDevice, st));
  int rc = solve_impl(s, precond, bd, xd, rtol, max_iter, hist_host, st, result);
  if (!rc) {
    cudaError_t e = cudaMemcpyAsync(x_host, xd, (size_t)s->n * sizeof(double), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      set_error(std::string("pcg_solve_host: copy back failed: ") + cudaGetErrorString(e));
      rc = B200SPAI_ERR_CUDA;
    }
  }
  dev_free(bd);
  dev_free(xd);
  return rc;
}

int b200spai_pcg_apply_precond(b200spai_pcg_t s, int precond, const double* r, double* z, void* stream) {
  B2_ARG(s && r && z, "pcg_apply_precond: NULL argument");
  B2_ARG(precond >= B200SPAI_PRECOND_EXT_SPAI && precond <= 3 && s->G, "pcg_apply_precond: SPAI preconditioner not set");
  cudaStream_t st = as_stream(stream);
  if (precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) B2_TRY(ensure_diag(s, st));
  int nl = 0;
  B2_TRY(enqueue_precond(s, precond, r, z, nullptr, st, &nl));
  { const long long _nl = (nl); s->launches += _nl; count_launches(_nl); }
  return 0;
}

int b200spai_pcg_profile_iterations(b200spai_pcg_t s, int precond, const double* b, double* x, int32_t iterations,
                                    float* kernel_ms5, void* stream) {
  B2_ARG(s && b && x && kernel_ms5 && iterations > 0, "pcg_profile_iterations: bad argument");
  B2_ARG(precond >= 0 && precond <= 3, "pcg_profile_iterations: unknown preconditioner");
  if (precond >= B200SPAI_PRECOND_EXT_SPAI) B2_ARG(s->G && s->GT, "pcg_profile_iterations: set_spai first");
  cudaStream_t st = as_stream(stream);
  if (precond == B200SPAI_PRECOND_DIAGONAL || precond == B200SPAI_PRECOND_EXT_SPAI_SCALED) B2_TRY(ensure_diag(s, st));
  B2_TRY(ensure_hist(s, (long long)iterations + 1));
  const bool spai = precond >= B200SPAI_PRECOND_EXT_SPAI;
  if (reinterpret_cast<uintptr_t>(x) & 15) x = s->xw;  // the vector kernels use 16-byte accesses
  PcgScalars init{};
  init.rtol = 0.0;  // never converges: exactly `iterations` updates are timed
  init.max_iter = iterations;
  *s->h_in = init;
  B2_CUDA(cudaMemcpyAsync(s->sc, s->h_in, sizeof(PcgScalars), cudaMemcpyHostToDevice, st));
  pcg_init_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(b, x, s->r, s->n, s->sc, s->partials, s->ticket, s->hist);
  B2_LAUNCH_CHECK();
  // one warm-up iteration builds the lazily created row partitions outside the timed events
  {
    int nl = 0;
    B2_TRY(enqueue_iteration(s, precond, x, st, &nl));
    *s->h_in = init;
    B2_CUDA(cudaMemcpyAsync(s->sc, s->h_in, sizeof(PcgScalars), cudaMemcpyHostToDevice, st));
    pcg_init_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(b, x, s->r, s->n, s->sc, s->partials, s->ticket, s->hist);
  }
  const int per = 6;  // event before each of the 5 kernels + one after the last
  std::vector<cudaEvent_t> ev((size_t)iterations * per);
  for (auto& e : ev) B2_CUDA(cudaEventCreate(&e));
  for (int it = 0; it < iterations; ++it) {
    cudaEvent_t* e = ev.data() + (size_t)it * per;
    int nl = 0;
    B2_CUDA(cudaEventRecord(e[0], st));
    if (spai) {
      const bool scaled = (precond == B200SPAI_PRECOND_EXT_SPAI_SCALED);
      SpmvArgs a1{};
      a1.x = s->r; a1.y = s->t; a1.d = scaled ? s->d : nullptr; a1.sc = s->sc;
      B2_TRY(spmv_launch(s->GT, a1, scaled ? SPMV_SCALE : SPMV_PLAIN, 0, st));
      B2_CUDA(cudaEventRecord(e[1], st));
      SpmvArgs a2{};
      a2.x = s->t; a2.y = s->z; a2.aux = s->r; a2.d = scaled ? s->d : nullptr; a2.eps = s->eps;
      a2.partials = s->partials; a2.ticket = s->ticket; a2.sc = s->sc; a2.final_op = FINAL_RHO;
      B2_TRY(spmv_launch(s->G, a2, SPMV_AXPY_DOT, 0, st));
    } else {
      B2_CUDA(cudaEventRecord(e[1], st));
      B2_TRY(enqueue_precond(s, precond, s->r, s->z, s->sc, st, &nl));
    }
    B2_CUDA(cudaEventRecord(e[2], st));
    pcg_update_p_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(s->z, s->p, s->n, s->sc);
    B2_CUDA(cudaEventRecord(e[3], st));
    SpmvArgs a{};
    a.x = s->p; a.aux = s->p; a.y = s->q; a.partials = s->partials; a.ticket = s->ticket; a.sc = s->sc; a.final_op = FINAL_PQ;
    B2_TRY(spmv_launch(s->A, a, SPMV_DOT, 0, st));
    B2_CUDA(cudaEventRecord(e[4], st));
    pcg_update_xr_kernel<<<vec_grid(s->n), VTHREADS, 0, st>>>(x, s->r, s->p, s->q, s->n, s->sc, s->partials, s->ticket,
                                                              s->hist);
    B2_CUDA(cudaEventRecord(e[5], st));
  }
  B2_LAUNCH_CHECK();
  B2_CUDA(cudaStreamSynchronize(st));
  for (int k = 0; k < 5; ++k) kernel_ms5[k] = 0.f;
  for (int it = 0; it < iterations; ++it)
    for (int k = 0; k < 5; ++k) {
      float ms = 0.f;
      B2_CUDA(cudaEventElapsedTime(&ms, ev[(size_t)it * per + k], ev[(size_t)it * per + k + 1]));
      kernel_ms5[k] += ms / iterations;
    }
  for (auto& e : ev) cudaEventDestroy(e);
  { const long long _nl = ((long long)iterations * (spai ? 5 : 4)); s->launches += _nl; count_launches(_nl); }
  return 0;
}

int b200spai_pcg_fused_profile(b200spai_pcg_t s, double* phase_ms5, int32_t* iterations, void* stream) {
  B2_ARG(s && phase_ms5, "pcg_fused_profile: NULL argument");
  B2_ARG(s->fprof, "pcg_fused_profile: no fused solve has run on this handle");
  unsigned long long h[6] = {0, 0, 0, 0, 0, 0};
  B2_CUDA(cudaMemcpyAsync(h, s->fprof, sizeof(h), cudaMemcpyDeviceToHost, as_stream(stream)));
  B2_CUDA(cudaStreamSynchronize(as_stream(stream)));
  for (int k = 0; k < 5; ++k) phase_ms5[k] = (double)h[k] * 1e-6;
  if (iterations) *iterations = (int32_t)h[5];
  return 0;
}

int b200spai_pcg_launch_count(b200spai_pcg_t s, int64_t* launches) {
  B2_ARG(s && launches, "pcg_launch_count: NULL argument");
  *launches = s->launches;
  return 0;
}

int b200spai_pcg_destroy(b200spai_pcg_t s) {
  if (!s) return 0;
  for (int i = 0; i < 4; ++i) {
    if (s->exec[i]) cudaGraphExecDestroy(s->exec[i]);
    if (s->graph[i]) cudaGraphDestroy(s->graph[i]);
  }
  if (s->GT) csr_free(s->GT);
  for (double* v : {s->r, s->z, s->p, s->q, s->t, s->d, s->bw, s->xw, s->partials, s->hist}) dev_free(v);
  dev_free(s->sc);
  dev_free(s->ticket);
  dev_free(s->fsync);
  dev_free(s->fprof);
  // the scaffold's streams and events are idle here: every solve synchronises before it returns
  scaffold_release(s);
  delete s;
  return 0;
}

}  // extern "C"
