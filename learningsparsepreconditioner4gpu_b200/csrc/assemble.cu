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
// PLAN (with FILL): also record where every stored value comes from (out_src: index into the flattened [E,b,b] values, -1
// for the unit diagonal a masked DOF receives) and append to zero_list every input entry that was dropped because its
// VALUE is exactly 0 while the mask keeps it -- the cached pattern is valid only while those stay 0.
template <bool FILL>
__global__ void expand_rows_masked_kernel(const long long* __restrict__ col, const float* __restrict__ vals,
                                          const float* __restrict__ mask, const int* __restrict__ brp, long long N, int b,
                                          int* __restrict__ counts, const int* __restrict__ rowptr, int* __restrict__ out_col,
                                          double* __restrict__ out_val, int* __restrict__ out_src = nullptr,
                                          int* __restrict__ zero_list = nullptr, unsigned long long* __restrict__ zero_count = nullptr,
                                          long long zero_cap = 0) {
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
        const bool killed = (mi == 0.0 || (double)mask[j] == 0.0);
        if (killed) v = 0.0;
        if (j > i && diag_pending) {  // diagonal was not stored: the identity creates it
          if (FILL) { out_col[w] = (int)i; out_val[w] = ident; if (out_src) out_src[w] = -1; ++w; }
          ++n;
          diag_pending = false;
        }
        bool from_ident = false;
        if (j == i) {
          v = v + ident;
          from_ident = killed && ident != 0.0;
          diag_pending = false;
        }
        if (v == 0.0) {  // scipy's sparse add drops exact zeros
          if (FILL && zero_list && !killed) {
            const unsigned long long pos = atomicAdd(zero_count, 1ull);
            if ((long long)pos < zero_cap) zero_list[pos] = (int)(((long long)e * b + bi) * b + bj);
          }
          continue;
        }
        if (FILL && out_src) out_src[w] = from_ident ? -1 : (int)(((long long)e * b + bi) * b + bj);
      } else if (FILL && out_src) {
        out_src[w] = (int)(((long long)e * b + bi) * b + bj);
      }
      if (FILL) { out_col[w] = (int)j; out_val[w] = v; ++w; }
      ++n;
    }
  }
  if (diag_pending) {
    if (FILL) { out_col[w] = (int)i; out_val[w] = ident; if (out_src) out_src[w] = -1; ++w; }
    ++n;
  }
  if (!FILL) counts[i] = n;
}

// ---- fixed-topology value refresh -----------------------------------------------------------------------------
__global__ void mask_binary_kernel(const float* __restrict__ mask, long long n, unsigned long long* __restrict__ flag) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && mask[i] != 0.0f && mask[i] != 1.0f) *flag = 1ull;
}

__global__ void refresh_from_edges_kernel(const int* __restrict__ src, const float* __restrict__ vals, long long nnz,
                                          double* __restrict__ out, float* __restrict__ out32, int* __restrict__ bad) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  int b = 0;
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
    const int s = src[k];
    const float f = s >= 0 ? vals[s] : 1.0f;
    b |= (s >= 0 && f == 0.0f);  // a stored entry became exactly 0: the reference would drop it from the pattern
    out[k] = (double)f;
    if (out32) out32[k] = f;
  }
  if (b) atomicOr(bad, 1);
}

__global__ void check_zero_list_kernel(const int* __restrict__ zl, const float* __restrict__ vals, long long n, int* __restrict__ bad) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  int b = 0;
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) b |= (vals[zl[k]] != 0.0f);
  if (b) atomicOr(bad, 1);
}

}  // namespace b200spai

using namespace b200spai;

static int edges_to_csr_impl(int64_t num_nodes, int64_t num_edges, int32_t block_size, const int64_t* edge_index_dev,
                            const float* values_dev, const float* mask_dev, void* stream, b200spai_csr_t* out, bool plan) {
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
  unsigned long long* zc = nullptr;
  long long zero_cap = 0;
  if (!rc && plan) {
    // every stored entry remembers its source; dropped value-zeros are listed (at most inputs - kept inputs of them)
    zero_cap = E * b * b - (long long)h_nnz + n + 1;
    rc = dev_alloc(&A->src, (size_t)h_nnz + 1);
    if (!rc) rc = dev_alloc(&A->zero_list, (size_t)zero_cap);
    if (!rc) rc = dev_alloc(&zc, 2);
    if (!rc && cudaMemsetAsync(zc, 0, 2 * sizeof(unsigned long long), st) != cudaSuccess) rc = B200SPAI_ERR_CUDA;
    if (!rc && mask_dev && n > 0) mask_binary_kernel<<<(unsigned)((n + tb - 1) / tb), tb, 0, st>>>(mask_dev, n, zc + 1);
    if (rc) {
      csr_free(A);
      A = nullptr;
    }
  }
  if (!rc) {
    cudaMemcpyAsync(A->rowptr, counts, ((size_t)n + 1) * sizeof(int), cudaMemcpyDeviceToDevice, st);
    if (n > 0)
      expand_rows_masked_kernel<true><<<(unsigned)((n + tb - 1) / tb), tb, 0, st>>>(col, values_dev, mask_dev, brp, N, b,
                                                                                   nullptr, A->rowptr, A->col, A->val,
                                                                                   plan ? A->src : nullptr, plan ? A->zero_list : nullptr,
                                                                                   zc, zero_cap);
    unsigned long long h_z[2] = {0, 0};
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && plan) e = cudaMemcpyAsync(h_z, zc, sizeof(h_z), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
      set_error(std::string("edges_to_csr: fill failed: ") + cudaGetErrorString(e));
      csr_free(A);
      A = nullptr;
      rc = B200SPAI_ERR_CUDA;
    } else if (plan) {
      A->n_zero = (long long)h_z[0] < zero_cap ? (long long)h_z[0] : zero_cap;
      A->n_src_values = E * b * b;
      A->src_kind = h_z[1] ? 0 : 1;  // a mask with entries other than 0 / 1 adds (1 - m) to live diagonals: no value-only refresh
      if (A->src_kind == 0) {
        dev_free(A->src);
        dev_free(A->zero_list);
        A->src = nullptr;
        A->zero_list = nullptr;
        A->n_zero = 0;
      }
    }
  }
  dev_free(zc);
  count_launches(plan ? 5 : 4);
  if (!rc) A->block_hint = b;
  if (!rc) *out = A;
  return rc;
}

extern "C" {

int b200spai_edges_to_csr(int64_t num_nodes, int64_t num_edges, int32_t block_size, const int64_t* edge_index_dev,
                          const float* values_dev, const float* mask_dev, void* stream, b200spai_csr_t* out) {
  return edges_to_csr_impl(num_nodes, num_edges, block_size, edge_index_dev, values_dev, mask_dev, stream, out, false);
}

int b200spai_edges_to_csr_plan(int64_t num_nodes, int64_t num_edges, int32_t block_size, const int64_t* edge_index_dev,
                               const float* values_dev, const float* mask_dev, void* stream, b200spai_csr_t* out) {
  return edges_to_csr_impl(num_nodes, num_edges, block_size, edge_index_dev, values_dev, mask_dev, stream, out, true);
}

int b200spai_csr_has_plan(b200spai_csr_t a, int32_t* kind) {
  B2_ARG(a && kind, "csr_has_plan: NULL argument");
  *kind = a->src_kind;
  return 0;
}

int b200spai_csr_refresh_from_edges(b200spai_csr_t a, const float* values_dev, int64_t num_values, void* stream, int32_t* ok) {
  B2_ARG(a && values_dev && ok, "csr_refresh_from_edges: NULL argument");
  B2_ARG(a->src_kind == 1 && a->src, "csr_refresh_from_edges: the matrix was not built by b200spai_edges_to_csr_plan (or its mask is not 0 / 1)");
  B2_ARG(num_values == a->n_src_values, "csr_refresh_from_edges: the value array does not have the planned E * b * b entries");
  cudaStream_t st = as_stream(stream);
  *ok = 0;
  int* bad = nullptr;
  B2_TRY(dev_alloc(&bad, 1));
  B2_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), st));
  const int tb = 256;
  const long long cap = (long long)sm_count() * 16;
  if (a->nnz > 0) {
    long long g = (a->nnz + tb * 4 - 1) / (tb * 4);
    refresh_from_edges_kernel<<<(unsigned)(g > cap ? cap : g), tb, 0, st>>>(a->src, values_dev, a->nnz, a->val,
                                                                           a->f32_state == 1 ? a->val32 : nullptr, bad);
  }
  if (a->n_zero > 0) {
    long long g = (a->n_zero + tb * 4 - 1) / (tb * 4);
    check_zero_list_kernel<<<(unsigned)(g > cap ? cap : g), tb, 0, st>>>(a->zero_list, values_dev, a->n_zero, bad);
  }
  B2_LAUNCH_CHECK();
  count_launches(a->n_zero > 0 ? 2 : 1);
  int h = 0;
  cudaError_t e = cudaMemcpyAsync(&h, bad, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  dev_free(bad);
  B2_CUDA(e);
  *ok = h ? 0 : 1;  // 0: an exact zero appeared or disappeared -- the pattern changed, rebuild the matrix
  if (*ok) B2_TRY(csr_values_changed(a, st));
  return 0;
}

}  // extern "C"
