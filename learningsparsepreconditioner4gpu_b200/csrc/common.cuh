// Shared helpers for libb200spai (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

#include "../../include/b200spai.h"

namespace b200spai {

void set_error(const std::string& msg);
int sm_count();  // cached SM count of the current device
void count_launches(long long n);  // process-wide tally of this library's kernel launches

#define B2_CUDA(call)                                                                           \
  do {                                                                                          \
    cudaError_t _e = (call);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      ::b200spai::set_error(std::string(#call) + " failed: " + cudaGetErrorString(_e) + " at " + \
                            __FILE__ + ":" + std::to_string(__LINE__));                         \
      return B200SPAI_ERR_CUDA;                                                                 \
    }                                                                                           \
  } while (0)

#define B2_ARG(cond, msg)                 \
  do {                                    \
    if (!(cond)) {                        \
      ::b200spai::set_error(msg);         \
      return B200SPAI_ERR_ARG;            \
    }                                     \
  } while (0)

#define B2_TRY(expr)          \
  do {                        \
    int _rc = (expr);         \
    if (_rc != 0) return _rc; \
  } while (0)

#define B2_LAUNCH_CHECK() B2_CUDA(cudaGetLastError())

// Device memory comes from a per-device caching pool (csr.cu): a solve step allocates and releases the same few
// multi-GB blocks every time, and cudaMalloc / cudaFree of those cost ~100 ms per step (cudaFree also synchronises
// the device).  The pool is STREAM-AWARE: every block is tagged with the stream of the C-ABI call that allocated it
// (as_stream() below records the calling thread's current stream) and a released block is handed only to a later
// allocation made for the SAME stream -- stream order then guarantees that the kernels still queued on the block when
// it was released (temporaries of csr_transpose / edges_to_csr are freed right after their kernels are enqueued) finish
// before the new owner's work starts.  Handles of different host threads / streams (batch.BatchSolver) therefore never
// recycle each other's in-flight blocks.  b200spai_pool_trim() returns cached blocks.
cudaError_t pool_malloc(void** p, size_t bytes);
void pool_free(void* p);
void pool_note_stream(cudaStream_t s);  // the stream the calling thread's allocations belong to from now on

template <typename T>
inline int dev_alloc(T** p, size_t count) {
  B2_CUDA(::b200spai::pool_malloc((void**)p, (count ? count : 1) * sizeof(T)));
  return 0;
}
inline void dev_free(void* p) { ::b200spai::pool_free(p); }

static inline cudaStream_t as_stream(void* s) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(s);
  pool_note_stream(st);
  return st;
}

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum in a fixed order (deterministic for a fixed blockDim). `red` holds >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` against a previous use
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  double t = (threadIdx.x < nw) ? red[threadIdx.x] : 0.0;
  if (warp == 0) t = warp_sum(t);
  return t;  // valid in warp 0 (all lanes)
}

// Grid-wide deterministic reduction: every CTA deposits its partial (valid in thread 0), the CTA
// that draws the last ticket sums the partials in index order and returns true in all its threads
// with the grand total in `total`.  `ticket` must be zero on entry and is reset by the last CTA.
// `red` holds >= 34 doubles of shared memory.
__device__ __forceinline__ bool grid_sum_last(double cta_partial, double* partials, unsigned* ticket, double* red,
                                              double& total) {
  __shared__ int s_is_last;
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = cta_partial;
    __threadfence();
    unsigned t = atomicAdd(ticket, 1u);
    s_is_last = (t == gridDim.x - 1) ? 1 : 0;
  }
  __syncthreads();
  if (!s_is_last) return false;
  __threadfence();
  double acc = 0.0;
  for (unsigned i = threadIdx.x; i < gridDim.x; i += blockDim.x) acc += __ldcg(partials + i);
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    red[33] = acc;
    *ticket = 0u;
  }
  __syncthreads();
  total = red[33];
  return true;
}

// ----- mbarrier / bulk-copy (TMA 1-D) PTX wrappers --------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy completing on an mbarrier; addresses and size multiples of 16 B.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar,
                                         uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(smem_dst)),
      "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace b200spai
