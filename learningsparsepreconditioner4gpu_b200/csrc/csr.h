// Device-resident CSR matrix + SpMV dispatch (internal header).
#pragma once
#include "bsr.cuh"

struct b200spai_csr_s {
  int n_rows = 0, n_cols = 0;
  long long nnz = 0;
  // device arrays, padded so that 16-byte bulk copies may over-read by up to 8 elements
  int* rowptr = nullptr;
  int* col = nullptr;
  double* val = nullptr;
  // fp32 copy of the values, present when every value is exactly representable in fp32 (built lazily with
  // the partition).  f32_state: 0 = not examined yet, 1 = val32 valid, -1 = not lossless / disabled.
  float* val32 = nullptr;
  int f32_state = 0;
  // 16-bit within-row column steps + first column of every row, present when every step fits 16 bits.
  // c16_state: 0 = not examined yet, 1 = col16/rowfirst valid, -1 = does not fit / disabled.
  unsigned short* col16 = nullptr;
  int* rowfirst = nullptr;
  int c16_state = 0;
  // streaming partition (built lazily per CAP: 1024, 2048, 4096, 8192): {r0, lo, r1, hi} per row block
  int4* blk[4] = {nullptr, nullptr, nullptr, nullptr};
  int nblk[4] = {0, 0, 0, 0};
  int max_row_len = -1;
  double mean_row_len = 0.0;
  // Native 3x3 block stream (bsr.cuh), built lazily for matrices assembled from [E,3,3] blocks (block_hint == 3) when it
  // is smaller than the scalar stream.  bsr_state: 0 = not examined yet, 1 = valid, -1 = does not apply / disabled.
  int block_hint = 1;
  int bsr_state = 0;
  int nb_rows = 0;
  long long nblocks = 0;
  int* browptr = nullptr;
  float* bval = nullptr;
  unsigned short* bstep = nullptr;
  int* bfirst = nullptr;
  int4* bblk = nullptr;
  int nbblk = 0;
  // Fixed-topology value refresh (b200spai_edges_to_csr_plan / b200spai_csr_transpose_plan): src[k] = where the value of
  // stored entry k comes from -- src_kind 1: index into the flattened [E,b,b] edge values (-1: the unit diagonal a
  // masked DOF receives); src_kind 2: position in the parent matrix's value array (this matrix is its transpose).
  // zero_list: edge-value indices that were dropped because the VALUE was exactly 0 (not because of the mask): the
  // cached pattern stays valid only while they remain 0.
  int* src = nullptr;
  int src_kind = 0;
  int* zero_list = nullptr;
  long long n_zero = 0;
  long long n_src_values = 0;  // src_kind 1: E * b * b

  b200spai::CsrView view() const { return b200spai::CsrView{n_rows, rowptr, col, val}; }
};

namespace b200spai {

// Allocate an empty matrix with padded arrays.
int csr_alloc(int n_rows, int n_cols, long long nnz, b200spai_csr_s** out);
void csr_free(b200spai_csr_s* a);

// kernel: 0 auto, 1 vector, 2 stream.  `mode` selects the epilogue.
int spmv_launch(b200spai_csr_s* A, const SpmvArgs& args, int mode, int kernel, cudaStream_t stream);
// Lazily built per-matrix pieces of the streaming kernel (row-block partition for CAP index `which`, fp32 value
// copy).  One-off; synchronises the stream.
int csr_ensure_partition(b200spai_csr_s* A, int which, cudaStream_t stream);
int csr_ensure_f32(b200spai_csr_s* A, cudaStream_t stream);
int csr_ensure_col16(b200spai_csr_s* A, cudaStream_t stream);
int csr_ensure_bsr3(b200spai_csr_s* A, cudaStream_t stream);
// The values of A changed in place (fixed-topology refresh): bring the derived value arrays (block stream) up to date.
int csr_values_changed(b200spai_csr_s* A, cudaStream_t stream);
// Contiguous row blocks {r0, rowptr[r0], r1, rowptr[r1]} balanced on rowptr[r] + weight*r (one-off, synchronises).
int build_row_blocks(const int* rowptr, int n_rows, long long nnz, long long target, int weight, int4** blk_out,
                     int* nblk_out, cudaStream_t stream);
// Resolve the kernel the auto policy picks (1 or 2).
int spmv_auto_kernel(const b200spai_csr_s* A);

}  // namespace b200spai
