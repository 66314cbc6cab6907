// CSR container, row-block partition, transpose and SpMV dispatch.
#include <cub/cub.cuh>

#include <atomic>
#include <cstdlib>
#include <map>
#include <tuple>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "csr.h"

namespace b200spai {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
const std::string& last_error() { return g_last_error; }

static std::atomic<long long> g_launches{0};
void count_launches(long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
long long total_launches() { return g_launches.load(std::memory_order_relaxed); }

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// ---- caching device-memory pool ------------------------------------------------------------------------------
namespace {
struct LiveBlock {
  size_t bytes;
  int dev;
  cudaStream_t stream;  // the stream of the call that allocated it: its kernels are ordered on this stream
};
struct Pool {
  std::mutex mu;
  std::unordered_map<void*, LiveBlock> live;
  // (device, bytes, stream) -> cached blocks.  A released block only goes back to an allocation for the SAME stream, so
  // work still queued on it when it was released is ordered before the new owner's (see common.cuh).
  std::map<std::tuple<int, size_t, cudaStream_t>, std::vector<void*>> free_blocks;
  size_t cached = 0, limit = 0;
};
thread_local cudaStream_t tls_stream = nullptr;
Pool& pool() {
  static Pool* p = new Pool();  // leaked on purpose: handles may be destroyed after static destructors ran
  return *p;
}
size_t pool_limit() {
  Pool& P = pool();
  if (P.limit == 0) {
    const char* e = getenv("B200SPAI_POOL_MAX_GB");
    const double gb = e ? atof(e) : 64.0;
    P.limit = (size_t)((gb < 0 ? 0 : gb) * (double)(1ull << 30)) + 1;
  }
  return P.limit;
}
// Size classes: multiples of 512 B up to 4 KB, above that 8 classes per power of two (<= 12.5 % slack), so that a
// stream of systems of DIFFERENT sizes (a batch of meshes) still hits the cache instead of cudaMalloc / cudaFree.
size_t round_size(size_t bytes) {
  bytes = (bytes + 511) & ~(size_t)511;
  if (bytes <= 4096) return bytes;
  int p = 63 - __builtin_clzll((unsigned long long)bytes);
  const size_t step = (size_t)1 << (p - 3);
  return (bytes + step - 1) & ~(step - 1);
}
}  // namespace

void pool_note_stream(cudaStream_t s) { tls_stream = s; }

cudaError_t pool_malloc(void** p, size_t bytes) {
  Pool& P = pool();
  bytes = round_size(bytes ? bytes : 1);
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const cudaStream_t st = tls_stream;
  {
    std::lock_guard<std::mutex> lock(P.mu);
    auto it = P.free_blocks.find(std::make_tuple(dev, bytes, st));
    if (it != P.free_blocks.end() && !it->second.empty()) {
      *p = it->second.back();
      it->second.pop_back();
      P.cached -= bytes;
      P.live[*p] = LiveBlock{bytes, dev, st};
      return cudaSuccess;
    }
  }
  e = cudaMalloc(p, bytes);
  if (e != cudaSuccess) {  // out of memory: give the cache back and try once more
    cudaGetLastError();
    b200spai_pool_trim();
    e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) return e;
  }
  std::lock_guard<std::mutex> lock(P.mu);
  P.live[*p] = LiveBlock{bytes, dev, st};
  return cudaSuccess;
}

void pool_free(void* p) {
  if (!p) return;
  Pool& P = pool();
  const size_t limit = pool_limit();
  {
    std::lock_guard<std::mutex> lock(P.mu);
    auto it = P.live.find(p);
    if (it != P.live.end()) {
      const LiveBlock blk = it->second;
      const size_t bytes = blk.bytes;
      P.live.erase(it);
      if (P.cached + bytes <= limit) {
        P.free_blocks[std::make_tuple(blk.dev, bytes, blk.stream)].push_back(p);
        P.cached += bytes;
        return;
      }
    }
  }
  cudaFree(p);  // over the cache limit, or not ours
}

static constexpr int PAD = 16;  // elements of slack behind every CSR array (bulk-copy over-read)

int csr_alloc(int n_rows, int n_cols, long long nnz, b200spai_csr_s** out) {
  B2_ARG(n_rows >= 0 && n_cols >= 0 && nnz >= 0, "csr_alloc: negative size");
  B2_ARG(nnz < 2147483647LL - 64, "csr_alloc: nnz must fit int32");
  auto* a = new b200spai_csr_s();
  a->n_rows = n_rows;
  a->n_cols = n_cols;
  a->nnz = nnz;
  int rc = dev_alloc(&a->rowptr, (size_t)n_rows + 1 + PAD);
  if (!rc) rc = dev_alloc(&a->col, (size_t)nnz + PAD);
  if (!rc) rc = dev_alloc(&a->val, (size_t)nnz + PAD);
  if (rc) {
    csr_free(a);
    return rc;
  }
  // define the padding (never used arithmetically, but keeps compute-sanitizer initcheck quiet)
  cudaMemset(a->rowptr + n_rows + 1, 0, PAD * sizeof(int));
  cudaMemset(a->col + nnz, 0, PAD * sizeof(int));
  cudaMemset(a->val + nnz, 0, PAD * sizeof(double));
  // these memsets run on the legacy default stream, which does NOT order with non-blocking streams: finish them
  // before the caller's stream touches the arrays
  cudaStreamSynchronize(cudaStreamLegacy);
  *out = a;
  return 0;
}

void csr_free(b200spai_csr_s* a) {
  if (!a) return;
  dev_free(a->rowptr);
  dev_free(a->col);
  dev_free(a->val);
  dev_free(a->val32);
  dev_free(a->col16);
  dev_free(a->rowfirst);
  dev_free(a->src);
  dev_free(a->zero_list);
  dev_free(a->browptr);
  dev_free(a->bval);
  dev_free(a->bstep);
  dev_free(a->bfirst);
  dev_free(a->bblk);
  dev_free(a->blk[0]);
  dev_free(a->blk[1]);
  dev_free(a->blk[2]);
  dev_free(a->blk[3]);
  delete a;
}

// ---- row-block partition -----------------------------------------------------------------------
// Block k starts at the first row r with key(r) = rowptr[r] + weight*r >= k*target, so a block holds
// < target entries (not counting its last row) and <= target/weight rows.
__global__ void partition_bounds_kernel(const int* __restrict__ rowptr, int n_rows, long long target, int weight,
                                        int nblk, int* __restrict__ bounds) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > nblk) return;
  const long long want = (long long)k * target;
  int lo = 0, hi = n_rows;
  while (lo < hi) {
    int mid = lo + ((hi - lo) >> 1);
    long long key = (long long)__ldg(rowptr + mid) + (long long)weight * mid;
    if (key >= want) hi = mid; else lo = mid + 1;
  }
  bounds[k] = (k == nblk) ? n_rows : lo;
}

__global__ void partition_pack_kernel(const int* __restrict__ rowptr, const int* __restrict__ bounds, int nblk,
                                      int4* __restrict__ blk) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nblk) return;
  int r0 = bounds[k], r1 = bounds[k + 1];
  blk[k] = make_int4(r0, __ldg(rowptr + r0), r1, __ldg(rowptr + r1));
}

int build_row_blocks(const int* rowptr, int n_rows, long long nnz, long long target, int weight, int4** blk_out,
                     int* nblk_out, cudaStream_t stream) {
  const long long total = nnz + (long long)weight * n_rows;
  const int nblk = (int)((total + target - 1) / target);
  *nblk_out = nblk;
  if (nblk == 0) return dev_alloc(blk_out, 1);
  int* bounds = nullptr;
  B2_TRY(dev_alloc(&bounds, (size_t)nblk + 1));
  B2_TRY(dev_alloc(blk_out, (size_t)nblk));
  partition_bounds_kernel<<<(nblk + 1 + 255) / 256, 256, 0, stream>>>(rowptr, n_rows, target, weight, nblk, bounds);
  partition_pack_kernel<<<(nblk + 255) / 256, 256, 0, stream>>>(rowptr, bounds, nblk, *blk_out);
  B2_LAUNCH_CHECK();
  count_launches(2);
  B2_CUDA(cudaStreamSynchronize(stream));
  dev_free(bounds);
  return 0;
}

static long long partition_target(int which) {
  return which == 0 ? StreamCfg<1024, double, int>::TARGET
                    : which == 1 ? StreamCfg<2048, double, int>::TARGET
                                 : which == 2 ? StreamCfg<4096, double, int>::TARGET : StreamCfg<8192, double, int>::TARGET;
}

int csr_ensure_partition(b200spai_csr_s* A, int which, cudaStream_t stream) {
  if (A->blk[which]) return 0;
  return build_row_blocks(A->rowptr, A->n_rows, A->nnz, partition_target(which), PART_W, &A->blk[which], &A->nblk[which], stream);
}

// ---- lossless fp32 value copy ---------------------------------------------------------------------
__global__ void narrow_values_kernel(const double* __restrict__ v, long long n, float* __restrict__ out, int* __restrict__ lossy) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  int bad = 0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double d = v[i];
    const float f = (float)d;
    out[i] = f;
    bad |= !((double)f == d);  // also true for NaN: keep the fp64 stream then
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(lossy, 1);
}

static bool f32_disabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200SPAI_NO_FP32");
    v = (e && atoi(e) != 0) ? 1 : 0;
  }
  return v == 1;
}

int csr_ensure_f32(b200spai_csr_s* A, cudaStream_t stream) {
  if (A->f32_state != 0) return 0;
  if (f32_disabled() || A->nnz == 0) {
    A->f32_state = -1;
    return 0;
  }
  int* lossy = nullptr;
  B2_TRY(dev_alloc(&lossy, 1));
  B2_CUDA(cudaMemsetAsync(lossy, 0, sizeof(int), stream));
  B2_TRY(dev_alloc(&A->val32, (size_t)A->nnz + PAD));
  B2_CUDA(cudaMemsetAsync(A->val32 + A->nnz, 0, PAD * sizeof(float), stream));
  long long ctas = (A->nnz + 1023) / 1024;
  if (ctas > (long long)sm_count() * 8) ctas = (long long)sm_count() * 8;
  narrow_values_kernel<<<(int)ctas, 256, 0, stream>>>(A->val, A->nnz, A->val32, lossy);
  B2_LAUNCH_CHECK();
  count_launches(1);
  int h = 0;
  B2_CUDA(cudaMemcpyAsync(&h, lossy, sizeof(int), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  dev_free(lossy);
  if (h) {
    dev_free(A->val32);
    A->val32 = nullptr;
    A->f32_state = -1;
  } else {
    A->f32_state = 1;
  }
  return 0;
}

// ---- lossless 16-bit column-step stream ----------------------------------------------------------------
__global__ void encode_col16_kernel(const int* __restrict__ rowptr, const int* __restrict__ col, int n_rows,
                                    unsigned short* __restrict__ col16, int* __restrict__ rowfirst, int* __restrict__ bad) {
  // one warp per row; lane k handles entries k, k+32, ... (each needs only its left neighbour)
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_rows) return;
  const int s = rowptr[w], e = rowptr[w + 1];
  if (lane == 0) rowfirst[w] = (s < e) ? col[s] : 0;
  int overflow = 0;
  for (int k = s + lane; k < e; k += 32) {
    const int d = (k == s) ? 0 : col[k] - col[k - 1];
    overflow |= (d < 0 || d > 65535);
    col16[k] = (unsigned short)d;
  }
  if (overflow) atomicOr(bad, 1);
}

static bool col16_disabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200SPAI_NO_COL16");
    v = (e && atoi(e) != 0) ? 1 : 0;
  }
  return v == 1;
}

int csr_ensure_col16(b200spai_csr_s* A, cudaStream_t stream) {
  if (A->c16_state != 0) return 0;
  if (col16_disabled() || A->nnz == 0) {
    A->c16_state = -1;
    return 0;
  }
  int* bad = nullptr;
  B2_TRY(dev_alloc(&bad, 1));
  B2_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), stream));
  B2_TRY(dev_alloc(&A->col16, (size_t)A->nnz + PAD));
  B2_TRY(dev_alloc(&A->rowfirst, (size_t)A->n_rows + 1 + PAD));
  B2_CUDA(cudaMemsetAsync(A->col16 + A->nnz, 0, PAD * sizeof(unsigned short), stream));
  B2_CUDA(cudaMemsetAsync(A->rowfirst + A->n_rows, 0, (1 + PAD) * sizeof(int), stream));
  const int tb = 256;
  encode_col16_kernel<<<(unsigned)(((long long)A->n_rows * 32 + tb - 1) / tb), tb, 0, stream>>>(A->rowptr, A->col, A->n_rows,
                                                                                            A->col16, A->rowfirst, bad);
  B2_LAUNCH_CHECK();
  count_launches(1);
  int h = 0;
  B2_CUDA(cudaMemcpyAsync(&h, bad, sizeof(int), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  dev_free(bad);
  if (h) {
    dev_free(A->col16);
    dev_free(A->rowfirst);
    A->col16 = nullptr;
    A->rowfirst = nullptr;
    A->c16_state = -1;
  } else {
    A->c16_state = 1;
  }
  return 0;
}

// ---- native 3x3 block stream (bsr.cuh) -------------------------------------------------------------------------
// One thread per block row walks its three scalar rows (sorted columns) as a 3-way merge over block columns.
// FILL = false: count the blocks; FILL = true: write the nine values (absent entries stay +0), the 16-bit block-column
// step and the first block column of the row.
template <bool FILL>
__global__ void bsr3_build_kernel(const int* __restrict__ rowptr, const int* __restrict__ col, const double* __restrict__ val,
                                  int nb_rows, int* __restrict__ counts, const int* __restrict__ browptr,
                                  float* __restrict__ bval, unsigned short* __restrict__ bstep, int* __restrict__ bfirst,
                                  int* __restrict__ bad) {
  const int R = blockIdx.x * blockDim.x + threadIdx.x;
  if (R >= nb_rows) return;
  int k[3], e[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    k[i] = rowptr[3 * R + i];
    e[i] = rowptr[3 * R + i + 1];
  }
  int n = 0, prev = 0;
  long long w = FILL ? browptr[R] : 0;
  while (true) {
    int cur = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < 3; ++i)
      if (k[i] < e[i]) cur = min(cur, col[k[i]] / 3);
    if (cur == 0x7fffffff) break;
    if (FILL) {
      float* out = bval + w * 9;
#pragma unroll
      for (int j = 0; j < 9; ++j) out[j] = 0.0f;
      const int step = (n == 0) ? 0 : cur - prev;
      if (step < 0 || step > 65535) atomicOr(bad, 1);
      bstep[w] = (unsigned short)step;
      if (n == 0) bfirst[R] = cur;
    }
#pragma unroll
    for (int i = 0; i < 3; ++i)
      while (k[i] < e[i] && col[k[i]] / 3 == cur) {
        if (FILL) bval[w * 9 + i * 3 + (col[k[i]] - 3 * cur)] = (float)val[k[i]];
        ++k[i];
      }
    prev = cur;
    ++n;
    ++w;
  }
  if (!FILL) counts[R] = n;
  else if (n == 0) bfirst[R] = 0;
}

static bool bsr_disabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200SPAI_NO_BSR");
    v = (e && atoi(e) != 0) ? 1 : 0;
  }
  return v == 1;
}

int csr_ensure_bsr3(b200spai_csr_s* A, cudaStream_t stream) {
  if (A->bsr_state != 0) return 0;
  A->bsr_state = -1;
  if (bsr_disabled() || A->block_hint != 3 || A->nnz == 0 || A->n_rows % 3 != 0 || A->n_cols % 3 != 0) return 0;
  B2_TRY(csr_ensure_f32(A, stream));
  if (A->f32_state != 1) return 0;  // the block stream carries fp32 values: only when that is lossless
  const int nb = A->n_rows / 3;
  const int tb = 128;
  int* counts = nullptr;
  B2_TRY(dev_alloc(&counts, (size_t)nb + 1 + PAD));
  B2_CUDA(cudaMemsetAsync(counts, 0, ((size_t)nb + 1 + PAD) * sizeof(int), stream));
  bsr3_build_kernel<false><<<(nb + tb - 1) / tb, tb, 0, stream>>>(A->rowptr, A->col, A->val, nb, counts, nullptr, nullptr,
                                                                 nullptr, nullptr, nullptr);
  B2_LAUNCH_CHECK();
  size_t tmp_bytes = 0;
  B2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, counts, nb + 1, stream));
  void* tmp = nullptr;
  B2_CUDA(pool_malloc(&tmp, tmp_bytes + 16));
  B2_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, counts, counts, nb + 1, stream));
  int h_nblocks = 0;
  B2_CUDA(cudaMemcpyAsync(&h_nblocks, counts + nb, sizeof(int), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  dev_free(tmp);
  // 38 bytes per block against 6 bytes per stored scalar entry: keep the block stream only when it is the smaller one
  if (38LL * h_nblocks >= 6LL * A->nnz) {
    dev_free(counts);
    return 0;
  }
  int* bad = nullptr;
  int rc = dev_alloc(&bad, 1);
  if (!rc) rc = dev_alloc(&A->bval, (size_t)h_nblocks * 9 + 64);
  if (!rc) rc = dev_alloc(&A->bstep, (size_t)h_nblocks + 64);
  if (!rc) rc = dev_alloc(&A->bfirst, (size_t)nb + 1 + PAD);
  if (rc) {
    dev_free(bad);
    dev_free(counts);
    return rc;
  }
  A->browptr = counts;
  A->nb_rows = nb;
  A->nblocks = h_nblocks;
  B2_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), stream));
  B2_CUDA(cudaMemsetAsync(A->bval + (size_t)h_nblocks * 9, 0, 64 * sizeof(float), stream));
  B2_CUDA(cudaMemsetAsync(A->bstep + h_nblocks, 0, 64 * sizeof(unsigned short), stream));
  B2_CUDA(cudaMemsetAsync(A->bfirst + nb, 0, (1 + PAD) * sizeof(int), stream));
  bsr3_build_kernel<true><<<(nb + tb - 1) / tb, tb, 0, stream>>>(A->rowptr, A->col, A->val, nb, nullptr, A->browptr, A->bval,
                                                                A->bstep, A->bfirst, bad);
  B2_LAUNCH_CHECK();
  count_launches(2);
  int h_bad = 0;
  B2_CUDA(cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, stream));
  B2_CUDA(cudaStreamSynchronize(stream));
  dev_free(bad);
  if (h_bad) return 0;  // a block-column step does not fit 16 bits: stay on the scalar streams (arrays freed with the matrix)
  B2_TRY(build_row_blocks(A->browptr, nb, h_nblocks, BSR_TARGET, BSR_W, &A->bblk, &A->nbblk, stream));
  A->bsr_state = 1;
  return 0;
}

int csr_values_changed(b200spai_csr_s* A, cudaStream_t stream) {
  if (A->bsr_state != 1) return 0;
  const int tb = 128;
  int* bad = nullptr;
  B2_TRY(dev_alloc(&bad, 1));
  B2_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), stream));
  bsr3_build_kernel<true><<<(A->nb_rows + tb - 1) / tb, tb, 0, stream>>>(A->rowptr, A->col, A->val, A->nb_rows, nullptr, A->browptr,
                                                                        A->bval, A->bstep, A->bfirst, bad);
  B2_LAUNCH_CHECK();
  count_launches(1);
  dev_free(bad);  // stream-ordered reuse (the pool hands it back to this stream only)
  return 0;
}

template <int MODE>
static int launch_bsr3(b200spai_csr_s* A, const SpmvArgs& args, cudaStream_t stream) {
  constexpr int STAGES = 2, NCW = 8, threads = (NCW + 1) * 32;
  const int smem = bsr_smem_bytes(STAGES);
  auto kern = spmv_bsr3_stream_kernel<STAGES, NCW, MODE>;
  static int ctas_per_sm = 0;  // per instantiation
  if (ctas_per_sm == 0) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int occ = 0;
    B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem));
    ctas_per_sm = occ > 0 ? occ : 1;
  }
  int grid = sm_count() * ctas_per_sm;
  if (grid > A->nbblk) grid = A->nbblk;
  if (grid < 1) grid = 1;
  BsrView B{A->nb_rows, A->browptr, A->bval, A->bstep, A->bfirst};
  kern<<<grid, threads, smem, stream>>>(B, A->bblk, A->nbblk, args);
  B2_LAUNCH_CHECK();
  return 0;
}

// ---- SpMV dispatch -------------------------------------------------------------------------------
static int stream_cfg_from_env() {
  static int cfg = -1;
  if (cfg < 0) {
    const char* e = getenv("B200SPAI_STREAM_CFG");
    cfg = e ? atoi(e) : 0;
    if (cfg < 0 || cfg > 4) cfg = 0;
  }
  return cfg;
}

int spmv_auto_kernel(const b200spai_csr_s* A) {
  static int forced = -1;
  if (forced < 0) {
    const char* e = getenv("B200SPAI_SPMV_KERNEL");
    forced = e ? atoi(e) : 0;
  }
  if (forced == 1 || forced == 2) return forced;
  // the streaming kernel needs enough row blocks to occupy the chip; tiny systems take the vector kernel
  return (A->nnz >= (1LL << 18)) ? 2 : 1;
}

template <int VEC, int MODE>
static int launch_vector(b200spai_csr_s* A, const SpmvArgs& args, cudaStream_t stream) {
  constexpr int RPW = 32 / VEC;
  const long long warps_needed = ((long long)A->n_rows + 2 * RPW - 1) / (2 * RPW);
  long long ctas = (warps_needed + 7) / 8;
  const long long cap = (long long)sm_count() * 8;
  if (ctas > cap) ctas = cap;
  if (ctas < 1) ctas = 1;
  spmv_vector_kernel<VEC, MODE><<<(int)ctas, 256, 0, stream>>>(A->view(), args);
  B2_LAUNCH_CHECK();
  return 0;
}

template <int MODE>
static int launch_vector_mode(b200spai_csr_s* A, const SpmvArgs& args, cudaStream_t stream) {
  const double m = A->n_rows ? (double)A->nnz / A->n_rows : 1.0;
  if (m <= 2.5) return launch_vector<2, MODE>(A, args, stream);
  if (m <= 5.0) return launch_vector<4, MODE>(A, args, stream);
  if (m <= 10.0) return launch_vector<8, MODE>(A, args, stream);
  if (m <= 24.0) return launch_vector<16, MODE>(A, args, stream);
  return launch_vector<32, MODE>(A, args, stream);
}

template <int CAP, int STAGES, int NCW, typename VT, typename CT, int MODE>
static int launch_stream_t(b200spai_csr_s* A, const VT* vals, const CT* cols, const SpmvArgs& args, cudaStream_t stream) {
  constexpr int which = (CAP == 1024) ? 0 : (CAP == 2048) ? 1 : (CAP == 4096) ? 2 : 3;
  constexpr int threads = (NCW + 1) * 32;
  const int smem = StreamCfg<CAP, VT, CT>::smem_bytes(STAGES);
  auto kern = spmv_stream_kernel<CAP, STAGES, NCW, VT, CT, MODE>;
  static int ctas_per_sm = 0;  // per instantiation
  if (ctas_per_sm == 0) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int occ = 0;
    B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem));
    ctas_per_sm = occ > 0 ? occ : 1;
  }
  int grid = sm_count() * ctas_per_sm;
  if (grid > A->nblk[which]) grid = A->nblk[which];
  if (grid < 1) grid = 1;
  kern<<<grid, threads, smem, stream>>>(A->view(), vals, cols, A->rowfirst, A->blk[which], A->nblk[which], args);
  B2_LAUNCH_CHECK();
  return 0;
}

template <int CAP, int STAGES, int NCW, int MODE>
static int launch_stream(b200spai_csr_s* A, const SpmvArgs& args, cudaStream_t stream) {
  constexpr int which = (CAP == 1024) ? 0 : (CAP == 2048) ? 1 : (CAP == 4096) ? 2 : 3;
  B2_TRY(csr_ensure_partition(A, which, stream));
  B2_TRY(csr_ensure_f32(A, stream));
  B2_TRY(csr_ensure_col16(A, stream));
  const bool f32 = A->f32_state == 1, c16 = A->c16_state == 1;
  if (f32 && c16) return launch_stream_t<CAP, STAGES, NCW, float, unsigned short, MODE>(A, A->val32, A->col16, args, stream);
  if (f32) return launch_stream_t<CAP, STAGES, NCW, float, int, MODE>(A, A->val32, A->col, args, stream);
  if (c16) return launch_stream_t<CAP, STAGES, NCW, double, unsigned short, MODE>(A, A->val, A->col16, args, stream);
  return launch_stream_t<CAP, STAGES, NCW, double, int, MODE>(A, A->val, A->col, args, stream);
}

template <int MODE>
static int launch_stream_mode(b200spai_csr_s* A, const SpmvArgs& args, cudaStream_t stream) {
  switch (stream_cfg_from_env()) {
    // development sweep (B200SPAI_STREAM_CFG = 1..4)
    case 1: return launch_stream<2048, 2, 4, MODE>(A, args, stream);
    case 2: return launch_stream<4096, 2, 6, MODE>(A, args, stream);
    case 3: return launch_stream<8192, 2, 8, MODE>(A, args, stream);
    case 4: return launch_stream<4096, 2, 8, MODE>(A, args, stream);
    default: break;
  }
  // measured on the 216^3 Poisson system (profiles/r01_spmv_sweep.txt): 4096-entry stages with 8 consumer warps
  // are the fastest once there are enough row blocks for every SM; mid-size systems keep the 2048-entry stages
  if (A->nnz >= (1LL << 22)) return launch_stream<4096, 2, 8, MODE>(A, args, stream);
  return launch_stream<2048, 2, 4, MODE>(A, args, stream);
}

template <int MODE>
static int spmv_launch_mode(b200spai_csr_s* A, const SpmvArgs& args, int kernel, cudaStream_t stream) {
  if (kernel == 0) kernel = spmv_auto_kernel(A);
  if (kernel == 2 && A->block_hint == 3) {  // assembled from 3x3 blocks: the native block stream when it is the smaller one
    B2_TRY(csr_ensure_bsr3(A, stream));
    if (A->bsr_state == 1) return launch_bsr3<MODE>(A, args, stream);
  }
  if (kernel == 2) return launch_stream_mode<MODE>(A, args, stream);
  return launch_vector_mode<MODE>(A, args, stream);
}

int spmv_launch(b200spai_csr_s* A, const SpmvArgs& args, int mode, int kernel, cudaStream_t stream) {
  switch (mode) {
    case SPMV_PLAIN: return spmv_launch_mode<SPMV_PLAIN>(A, args, kernel, stream);
    case SPMV_SCALE: return spmv_launch_mode<SPMV_SCALE>(A, args, kernel, stream);
    case SPMV_AXPY_DOT: return spmv_launch_mode<SPMV_AXPY_DOT>(A, args, kernel, stream);
    case SPMV_DOT: return spmv_launch_mode<SPMV_DOT>(A, args, kernel, stream);
  }
  set_error("spmv_launch: bad mode");
  return B200SPAI_ERR_ARG;
}

// ---- transpose ------------------------------------------------------------------------------------
__global__ void expand_rows_kernel(const int* __restrict__ rowptr, int n_rows, int* __restrict__ rowidx) {
  // one warp per row
  const long long w = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (w >= n_rows) return;
  const int s = rowptr[w], e = rowptr[w + 1];
  for (int k = s + lane; k < e; k += 32) rowidx[k] = (int)w;
}

__global__ void iota_kernel(int* p, long long n) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = (int)i;
}

__global__ void count_cols_kernel(const int* __restrict__ col, long long nnz, int* __restrict__ cnt) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nnz) atomicAdd(cnt + col[i] + 1, 1);  // integer counts: order-independent
}

__global__ void gather_transposed_kernel(const int* __restrict__ perm, const int* __restrict__ rowidx,
                                         const double* __restrict__ val, long long nnz, int* __restrict__ colT,
                                         double* __restrict__ valT) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nnz) return;
  const int src = perm[i];
  colT[i] = rowidx[src];
  valT[i] = val[src];
}

// value-only refresh of a matrix whose entries are a permutation of another value array (transpose plan)
__global__ void refresh_permuted_kernel(const int* __restrict__ src, const double* __restrict__ val, long long nnz,
                                        double* __restrict__ out, float* __restrict__ out32) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += stride) {
    const double v = val[src[i]];
    out[i] = v;
    if (out32) out32[i] = (float)v;
  }
}

static int csr_transpose_impl(b200spai_csr_s* A, cudaStream_t stream, b200spai_csr_s** out, bool keep_plan = false) {
  b200spai_csr_s* T = nullptr;
  B2_TRY(csr_alloc(A->n_cols, A->n_rows, A->nnz, &T));
  T->block_hint = A->block_hint;
  const long long nnz = A->nnz;
  const int tb = 256;
  B2_CUDA(cudaMemsetAsync(T->rowptr, 0, ((size_t)A->n_cols + 1) * sizeof(int), stream));
  if (nnz > 0) {
    int *rowidx = nullptr, *iota = nullptr, *keys_out = nullptr, *perm = nullptr;
    B2_TRY(dev_alloc(&rowidx, nnz));
    B2_TRY(dev_alloc(&iota, nnz));
    B2_TRY(dev_alloc(&keys_out, nnz));
    B2_TRY(dev_alloc(&perm, nnz));
    const long long warps = A->n_rows;
    expand_rows_kernel<<<(unsigned)((warps * 32 + tb - 1) / tb), tb, 0, stream>>>(A->rowptr, A->n_rows, rowidx);
    iota_kernel<<<(unsigned)((nnz + tb - 1) / tb), tb, 0, stream>>>(iota, nnz);
    count_cols_kernel<<<(unsigned)((nnz + tb - 1) / tb), tb, 0, stream>>>(A->col, nnz, T->rowptr);
    B2_LAUNCH_CHECK();
    // stable LSD radix sort of (column, position): within a column, ascending original position = ascending row
    int end_bit = 1;
    while (end_bit < 31 && (1LL << end_bit) < (long long)A->n_cols) ++end_bit;
    size_t tmp_bytes = 0;
    B2_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, A->col, keys_out, iota, perm, (int)nnz, 0, end_bit, stream));
    size_t scan_bytes = 0;
    B2_CUDA(cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, T->rowptr, T->rowptr, A->n_cols + 1, stream));
    void* tmp = nullptr;
    B2_CUDA(pool_malloc(&tmp, (tmp_bytes > scan_bytes ? tmp_bytes : scan_bytes) + 16));
    B2_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, A->col, keys_out, iota, perm, (int)nnz, 0, end_bit, stream));
    B2_CUDA(cub::DeviceScan::InclusiveSum(tmp, scan_bytes, T->rowptr, T->rowptr, A->n_cols + 1, stream));
    gather_transposed_kernel<<<(unsigned)((nnz + tb - 1) / tb), tb, 0, stream>>>(perm, rowidx, A->val, nnz, T->col, T->val);
    B2_LAUNCH_CHECK();
    count_launches(4);  // + the cub sort / scan passes, which are library kernels and not counted
    B2_CUDA(cudaStreamSynchronize(stream));
    dev_free(tmp);
    dev_free(rowidx);
    dev_free(iota);
    dev_free(keys_out);
    if (keep_plan) {
      T->src = perm;
      T->src_kind = 2;
    } else {
      dev_free(perm);
    }
  }
  *out = T;
  return 0;
}

__global__ void diagonal_kernel(CsrView A, double* __restrict__ diag) {
  int row = blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= A.n_rows) return;
  double d = 0.0;
  for (int k = A.rowptr[row]; k < A.rowptr[row + 1]; ++k)
    if (A.col[k] == row) d += A.val[k];
  diag[row] = d;
}

}  // namespace b200spai

using namespace b200spai;

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

const char* b200spai_last_error(void) { return last_error().c_str(); }
int b200spai_version(void) { return 102; }

int b200spai_csr_stream_info(b200spai_csr_t a, int32_t* value_bytes, int32_t* index_bytes, void* stream) {
  B2_ARG(a, "csr_stream_info: NULL handle");
  B2_TRY(csr_ensure_f32(a, as_stream(stream)));
  B2_TRY(csr_ensure_col16(a, as_stream(stream)));
  if (value_bytes) *value_bytes = (a->f32_state == 1) ? 4 : 8;
  if (index_bytes) *index_bytes = (a->c16_state == 1) ? 2 : 4;
  return 0;
}

int b200spai_csr_bsr_info(b200spai_csr_t a, void* stream, int32_t* block_size, int64_t* nblocks) {
  B2_ARG(a, "csr_bsr_info: NULL handle");
  if (a->block_hint == 3) B2_TRY(csr_ensure_bsr3(a, as_stream(stream)));
  if (block_size) *block_size = (a->bsr_state == 1) ? 3 : 1;
  if (nblocks) *nblocks = (a->bsr_state == 1) ? a->nblocks : 0;
  return 0;
}

int b200spai_pool_trim(void) {
  std::vector<void*> blocks;
  {
    Pool& P = pool();
    std::lock_guard<std::mutex> lock(P.mu);
    for (auto& kv : P.free_blocks) {
      blocks.insert(blocks.end(), kv.second.begin(), kv.second.end());
      kv.second.clear();
    }
    P.cached = 0;
  }
  cudaDeviceSynchronize();  // a cached block may still be read by work enqueued before its release
  for (void* b : blocks) cudaFree(b);
  return 0;
}

int b200spai_pool_cached_bytes(int64_t* bytes) {
  B2_ARG(bytes, "pool_cached_bytes: NULL argument");
  Pool& P = pool();
  std::lock_guard<std::mutex> lock(P.mu);
  *bytes = (int64_t)P.cached;
  return 0;
}

/* development hook (not in the public header): cycle counters of the streaming SpMV when built with -DB200SPAI_PROBE */
int b200spai_debug_probe(unsigned long long* out8, int reset) {
#ifdef B200SPAI_PROBE
  if (out8) B2_CUDA(cudaMemcpyFromSymbol(out8, g_probe, 8 * sizeof(unsigned long long)));
  if (reset) {
    unsigned long long z[8] = {0};
    B2_CUDA(cudaMemcpyToSymbol(g_probe, z, sizeof(z)));
  }
  return 0;
#else
  (void)out8;
  (void)reset;
  set_error("built without B200SPAI_PROBE");
  return B200SPAI_ERR_ARG;
#endif
}

int b200spai_total_launches(int64_t* launches) {
  B2_ARG(launches, "total_launches: NULL argument");
  *launches = total_launches();
  return 0;
}

int b200spai_device_info(char* name128, int* sm, int* cc) {
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp p;
  B2_CUDA(cudaGetDeviceProperties(&p, dev));
  if (name128) snprintf(name128, 128, "%s", p.name);
  if (sm) *sm = p.multiProcessorCount;
  if (cc) *cc = p.major * 10 + p.minor;
  return 0;
}

static int csr_create_common(int32_t n_rows, int32_t n_cols, int64_t nnz, const int32_t* indptr, const int32_t* indices,
                             const double* data, cudaMemcpyKind kind, void* stream, b200spai_csr_t* out) {
  B2_ARG(out != nullptr, "csr_create: out is NULL");
  B2_ARG(indptr && (nnz == 0 || (indices && data)), "csr_create: NULL array");
  b200spai_csr_s* a = nullptr;
  B2_TRY(csr_alloc(n_rows, n_cols, nnz, &a));
  cudaStream_t st = as_stream(stream);
  B2_CUDA(cudaMemcpyAsync(a->rowptr, indptr, ((size_t)n_rows + 1) * sizeof(int), kind, st));
  if (nnz > 0) {
    B2_CUDA(cudaMemcpyAsync(a->col, indices, (size_t)nnz * sizeof(int), kind, st));
    B2_CUDA(cudaMemcpyAsync(a->val, data, (size_t)nnz * sizeof(double), kind, st));
  }
  if (kind == cudaMemcpyHostToDevice) B2_CUDA(cudaStreamSynchronize(st));  // caller may free host arrays
  *out = a;
  return 0;
}

int b200spai_csr_create_host(int32_t n_rows, int32_t n_cols, int64_t nnz, const int32_t* indptr, const int32_t* indices,
                             const double* data, void* stream, b200spai_csr_t* out) {
  return csr_create_common(n_rows, n_cols, nnz, indptr, indices, data, cudaMemcpyHostToDevice, stream, out);
}

int b200spai_csr_create_dev(int32_t n_rows, int32_t n_cols, int64_t nnz, const int32_t* indptr, const int32_t* indices,
                            const double* data, void* stream, b200spai_csr_t* out) {
  return csr_create_common(n_rows, n_cols, nnz, indptr, indices, data, cudaMemcpyDeviceToDevice, stream, out);
}

int b200spai_csr_transpose(b200spai_csr_t a, void* stream, b200spai_csr_t* out) {
  B2_ARG(a && out, "csr_transpose: NULL handle");
  return csr_transpose_impl(a, as_stream(stream), out);
}

int b200spai_csr_transpose_plan(b200spai_csr_t a, void* stream, b200spai_csr_t* out) {
  B2_ARG(a && out, "csr_transpose_plan: NULL handle");
  return csr_transpose_impl(a, as_stream(stream), out, true);
}

int b200spai_csr_refresh_transposed(b200spai_csr_t t, b200spai_csr_t a, void* stream) {
  B2_ARG(t && a, "csr_refresh_transposed: NULL handle");
  B2_ARG(t->nnz == 0 || (t->src_kind == 2 && t->src), "csr_refresh_transposed: t was not built by b200spai_csr_transpose_plan");
  B2_ARG(t->nnz == a->nnz && t->n_rows == a->n_cols && t->n_cols == a->n_rows, "csr_refresh_transposed: t is not the transpose of a");
  if (t->nnz == 0) return 0;
  long long g = (t->nnz + 1023) / 1024;
  const long long cap = (long long)sm_count() * 16;
  refresh_permuted_kernel<<<(unsigned)(g > cap ? cap : g), 256, 0, as_stream(stream)>>>(t->src, a->val, t->nnz, t->val,
                                                                                       t->f32_state == 1 ? t->val32 : nullptr);
  B2_LAUNCH_CHECK();
  count_launches(1);
  return csr_values_changed(t, as_stream(stream));
}

int b200spai_csr_shape(b200spai_csr_t a, int32_t* n_rows, int32_t* n_cols, int64_t* nnz) {
  B2_ARG(a, "csr_shape: NULL handle");
  if (n_rows) *n_rows = a->n_rows;
  if (n_cols) *n_cols = a->n_cols;
  if (nnz) *nnz = a->nnz;
  return 0;
}

int b200spai_csr_download(b200spai_csr_t a, int32_t* indptr, int32_t* indices, double* data, void* stream) {
  B2_ARG(a, "csr_download: NULL handle");
  cudaStream_t st = as_stream(stream);
  if (indptr) B2_CUDA(cudaMemcpyAsync(indptr, a->rowptr, ((size_t)a->n_rows + 1) * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (indices && a->nnz) B2_CUDA(cudaMemcpyAsync(indices, a->col, (size_t)a->nnz * sizeof(int), cudaMemcpyDeviceToHost, st));
  if (data && a->nnz) B2_CUDA(cudaMemcpyAsync(data, a->val, (size_t)a->nnz * sizeof(double), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  return 0;
}

int b200spai_csr_device_ptrs(b200spai_csr_t a, const int32_t** indptr, const int32_t** indices, const double** data) {
  B2_ARG(a, "csr_device_ptrs: NULL handle");
  if (indptr) *indptr = a->rowptr;
  if (indices) *indices = a->col;
  if (data) *data = a->val;
  return 0;
}

int b200spai_csr_destroy(b200spai_csr_t a) {
  csr_free(a);
  return 0;
}

int b200spai_spmv(b200spai_csr_t a, const double* x, double* y, int kernel, void* stream) {
  B2_ARG(a && x && y, "spmv: NULL argument");
  B2_ARG(kernel >= 0 && kernel <= 2, "spmv: kernel must be 0, 1 or 2");
  if (a->n_rows == 0) return 0;
  SpmvArgs args{};
  args.x = x;
  args.y = y;
  count_launches(1);
  return spmv_launch(a, args, SPMV_PLAIN, kernel, as_stream(stream));
}

int b200spai_csr_diagonal(b200spai_csr_t a, double* diag, void* stream) {
  B2_ARG(a && diag, "csr_diagonal: NULL argument");
  if (a->n_rows == 0) return 0;
  diagonal_kernel<<<(a->n_rows + 255) / 256, 256, 0, as_stream(stream)>>>(a->view(), diag);
  B2_LAUNCH_CHECK();
  count_launches(1);
  return 0;
}

}  // extern "C"
