// Edge list (+ optional Dirichlet mask) -> fp64 CSR on the device.
//
// Device restatement of to_csr_cpu (neural_cg/utils/validate.py:22-51): [E,b,b] fp32 blocks at
// (edge_index[0], edge_index[1]) are expanded to scalar entries (row-major inside a block); with a mask,
// every entry whose row or column DOF has mask == 0 is zeroed and (1 - mask[i]) is added on the
// diagonal (apply_dbc_masking, neural_cg/data.py:159-170) -- which, because the reference does it with a
// scipy sparse add, also drops every entry that is exactly 0.0 afterwards, and creates the diagonal
// entry when it was not stored.  Without a mask nothing is dropped.  Output indices ascend inside rows.
//
// Requires the reference's own edge ordering (row-major, ascending columns, no duplicates --
// neural_cg/data.py:442-450); anything else is rejected with B200SPAI_ERR_UNSORTED.
#include <cub/cub.cuh>

#include "csr.h"

namespace b200spai {

__global__ void check_sorted_kernel(const long long* __restrict__ row, const long long* __restrict__ col, long long E,
                                    long long N, int* __restrict__ bad) {
  long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const long long r = row[e], c = col[e];
  bool ok = (r >= 0 && r < N && c >= 0 && c < N);
  if (e + 1 < E) {
    const long long r2 = row[e + 1], c2 = col[e + 1];
    ok = ok && (r2 > r || (r2 == r && c2 > c));
  }
  if (!ok) atomicOr(bad, 1);
}

// block row pointer: first edge whose row >= r
__global__ void block_rowptr_kernel(const long long* __restrict__ row, long long E, long long N, int* __restrict__ brp) {
  long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > N) return;
  long long lo = 0, hi = E;
  while (lo < hi) {
    long long mid = lo + ((hi - lo) >> 1);
    if (row[mid] >= r) hi = mid; else lo = mid + 1;
  }
  brp[r] = (int)lo;
}

// One thread per scalar row. FILL = false: count entries; FILL = true: write them.
template <bool FILL>
__global__ void expand_rows_masked_kernel(const long long* __restrict__ col, const float* __restrict__ vals,
                                          const float* __restrict__ mask, const int* __restrict__ brp, long long N, int b,
                                          int* __restrict__ counts, const int* __restrict__ rowptr, int* __restrict__ out_col,
                                          double* __restrict__ out_val) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // scalar row
  if (i >= N * b) return;
  const long long r = i / b;
  const int bi = (int)(i - r * b);
  const bool masked = (mask != nullptr);
  const double mi = masked ? (double)mask[i] : 1.0;
  const double ident = masked ? 1.0 - mi : 0.0;
  bool diag_pending = masked && (ident != 0.0);
  int n = 0;
  int w = FILL ? rowptr[i] : 0;
  for (int e = brp[r]; e < brp[r + 1]; ++e) {
    const long long cb = col[e];
    if (cb < 0 || cb >= N) continue;  // out-of-range column: check_sorted_kernel has flagged it and the call fails; never index mask with it
    const float* blk = vals + ((long long)e * b + bi) * b;
    for (int bj = 0; bj < b; ++bj) {
      const long long j = cb * b + bj;
      double v = (double)blk[bj];
      if (masked) {
        if (mi == 0.0 || (double)mask[j] == 0.0) v = 0.0;
        if (j > i && diag_pending) {  // diagonal was not stored: the identity creates it
          if (FILL) { out_col[w] = (int)i; out_val[w] = ident; ++w; }
          ++n;
          diag_pending = false;
        }
        if (j == i) {
          v = v + ident;
          diag_pending = false;
        }
        if (v == 0.0) continue;  // scipy's sparse add drops exact zeros
      }
      if (FILL) { out_col[w] = (int)j; out_val[w] = v; ++w; }
      ++n;
    }
  }
  if (diag_pending) {
    if (FILL) { out_col[w] = (int)i; out_val[w] = ident; ++w; }
    ++n;
  }
  if (!FILL) counts[i] = n;
}

}  // namespace b200spai

using namespace b200spai;

extern "C" int b200spai_edges_to_csr(int64_t num_nodes, int64_t num_edges, int32_t block_size,
                                     const int64_t* edge_index_dev, const float* values_dev, const float* mask_dev,
                                     void* stream, b200spai_csr_t* out) {
  B2_ARG(out && edge_index_dev && values_dev, "edges_to_csr: NULL argument");
  B2_ARG(block_size >= 1 && block_size <= 8, "edges_to_csr: block_size must be in [1,8]");
  B2_ARG(num_nodes >= 0 && num_edges >= 0, "edges_to_csr: negative size");
  const long long N = num_nodes, E = num_edges;
  const int b = block_size;
  const long long n = N * b;
  B2_ARG(n < 2147483647LL && E * b * b < 2147483647LL - 64, "edges_to_csr: sizes must fit int32");
  cudaStream_t st = as_stream(stream);
  const long long* row = reinterpret_cast<const long long*>(edge_index_dev);
  const long long* col = row + E;
  const int tb = 256;

  // temporaries are released on every return path (the B2_TRY / B2_CUDA macros return early)
  struct Temps {
    int *bad = nullptr, *brp = nullptr, *counts = nullptr;
    void* tmp = nullptr;
    ~Temps() {
      dev_free(tmp);
      dev_free(bad);
      dev_free(brp);
      dev_free(counts);
    }
  } T;
  int*& bad = T.bad;
  B2_TRY(dev_alloc(&bad, 1));
  B2_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), st));
  if (E > 0) check_sorted_kernel<<<(unsigned)((E + tb - 1) / tb), tb, 0, st>>>(row, col, E, N, bad);
  int*& brp = T.brp;
  B2_TRY(dev_alloc(&brp, (size_t)N + 1));
  block_rowptr_kernel<<<(unsigned)((N + 1 + tb - 1) / tb), tb, 0, st>>>(row, E, N, brp);
  int*& counts = T.counts;  // becomes the row pointer after the scan (n + 1 entries)
  B2_TRY(dev_alloc(&counts, (size_t)n + 1));
  B2_CUDA(cudaMemsetAsync(counts, 0, ((size_t)n + 1) * sizeof(int), st));
  if (n > 0)
    expand_rows_masked_kernel<false><<<(unsigned)((n + tb - 1) / tb), tb, 0, st>>>(col, values_dev, mask_dev, brp, N, b,
                                                                                  counts, nullptr, nullptr, nullptr);
  B2_LAUNCH_CHECK();
  size_t tmp_bytes = 0;
  B2_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, counts, counts, (int)(n + 1), st));
  void*& tmp = T.tmp;
  B2_CUDA(pool_malloc(&tmp, tmp_bytes + 16));
  B2_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, counts, counts, (int)(n + 1), st));
  int h_bad = 0, h_nnz = 0;
  B2_CUDA(cudaMemcpyAsync(&h_bad, bad, sizeof(int), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaMemcpyAsync(&h_nnz, counts + n, sizeof(int), cudaMemcpyDeviceToHost, st));
  B2_CUDA(cudaStreamSynchronize(st));
  int rc = 0;
  b200spai_csr_s* A = nullptr;
  if (h_bad) {
    set_error("edges_to_csr: edge_index must be in range, row-major, column-ascending and duplicate-free");
    rc = B200SPAI_ERR_UNSORTED;
  }
  if (!rc) rc = csr_alloc((int)n, (int)n, h_nnz, &A);
  if (!rc) {
    cudaMemcpyAsync(A->rowptr, counts, ((size_t)n + 1) * sizeof(int), cudaMemcpyDeviceToDevice, st);
    if (n > 0)
      expand_rows_masked_kernel<true><<<(unsigned)((n + tb - 1) / tb), tb, 0, st>>>(col, values_dev, mask_dev, brp, N, b,
                                                                                   nullptr, A->rowptr, A->col, A->val);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      set_error(std::string("edges_to_csr: fill failed: ") + cudaGetErrorString(e));
      csr_free(A);
      rc = B200SPAI_ERR_CUDA;
    }
  }
  count_launches(4);
  if (!rc) *out = A;
  return rc;
}
