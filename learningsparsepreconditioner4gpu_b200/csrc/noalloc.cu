// Allocation-free variant of the SPAI preconditioner apply  z = G (G^T r) + eps r  (b200spai_pcg_apply_precond, "ext_spai").
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The existing entry point keeps the intermediate t = G^T r in a buffer its solver handle allocated.  Here t lives in a
// file-scope `static __device__` array of B200SPAI_NOALLOC_MAX_ROWS doubles (8 MiB; every GPU that loads the library holds
// it until the process exits).  The entry point launches two kernels and allocates nothing: no cudaMalloc / cudaFree, no
// memory-pool call, no lazily built stream or partition of the matrices (its own row-per-thread CSR kernels read the
// canonical rowptr / col / val arrays directly), no torch call.
//
//   * `s_scratch_t` is, in HOST code, only a stand-in for the symbol -- not a device address.  The device address is
//     fetched once per device, with that device current, by cudaGetSymbolAddress (checked) and passed to the kernels.
//   * every call shares the one array and finds the previous call's contents in it.  Nothing has to start at zero: kernel 1
//     overwrites entries [0, n) before kernel 2 reads exactly those.  Calls that OVERLAP on the GPU (two streams, two host
//     threads) race on the array -- serialise them.
//   * row sums are formed sequentially with a rounded product and a rounded add, in ascending column order, like every
//     other SpMV of the library (scipy's csr_matvec bits).
#include "csr.h"

namespace b200spai {

static __device__ __align__(256) double s_scratch_t[B200SPAI_NOALLOC_MAX_ROWS];

constexpr int NA_MAX_DEVICES = 64;
static double* g_scratch_addr[NA_MAX_DEVICES] = {};  // per device: the DEVICE address of s_scratch_t (0 = not fetched yet)

static int scratch_address(double** out) {
  int dev = -1;
  B2_CUDA(cudaGetDevice(&dev));
  B2_ARG(dev >= 0 && dev < NA_MAX_DEVICES, "spai_apply_noalloc: device ordinal out of range");
  if (!g_scratch_addr[dev]) {
    double* p = nullptr;
    B2_CUDA(cudaGetSymbolAddress((void**)&p, s_scratch_t));
    B2_ARG(p != nullptr, "spai_apply_noalloc: cudaGetSymbolAddress returned NULL");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

// y[i] = sum_k val[k] x[col[k]]  (+ eps * aux[i] when aux != NULL), thread per row
__global__ void __launch_bounds__(256) na_spmv_kernel(int n, const int* __restrict__ rowptr, const int* __restrict__ col,
                                                      const double* __restrict__ val, const double* __restrict__ x,
                                                      const double* __restrict__ aux, double eps, double* __restrict__ y) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double acc = 0.0;
  for (int k = rowptr[i]; k < rowptr[i + 1]; ++k) acc = __dadd_rn(acc, __dmul_rn(val[k], x[col[k]]));
  y[i] = aux ? __dadd_rn(acc, __dmul_rn(eps, aux[i])) : acc;
}

}  // namespace b200spai

using namespace b200spai;

extern "C" int b200spai_spai_apply_noalloc(b200spai_csr_t G, b200spai_csr_t GT, double epsilon, const double* r_dev,
                                           double* z_dev, void* stream) {
  B2_ARG(G && GT && r_dev && z_dev, "spai_apply_noalloc: NULL argument");
  B2_ARG(G->n_rows == G->n_cols && GT->n_rows == G->n_cols && GT->n_cols == G->n_rows && GT->nnz == G->nnz,
         "spai_apply_noalloc: GT must be the transpose of the square matrix G");
  const int n = G->n_rows;
  if (n > B200SPAI_NOALLOC_MAX_ROWS) {
    set_error("spai_apply_noalloc: " + std::to_string(n) + " rows exceed the static scratch of B200SPAI_NOALLOC_MAX_ROWS = " +
              std::to_string(B200SPAI_NOALLOC_MAX_ROWS) + " rows; use b200spai_pcg_apply_precond");
    return B200SPAI_ERR_ARG;
  }
  if (n == 0) return 0;
  double* t = nullptr;
  B2_TRY(scratch_address(&t));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int grid = (n + 255) / 256;
  na_spmv_kernel<<<grid, 256, 0, st>>>(n, GT->rowptr, GT->col, GT->val, r_dev, nullptr, 0.0, t);
  na_spmv_kernel<<<grid, 256, 0, st>>>(n, G->rowptr, G->col, G->val, t, r_dev, epsilon, z_dev);
  B2_LAUNCH_CHECK();
  count_launches(2);
  return 0;
}
