// TEST INFRASTRUCTURE (not product): a minimal host stand-in for the CUDA execution model, enough to compile the
// backward kernels of csrc/gnn_bwd.cuh and their driver (csrc/gnn_train.cuh) with g++ and run them on the CPU.
//
//   * one host thread per CUDA thread of a CTA; CTAs of a grid run one after the other;
//   * __syncthreads() / __syncwarp() are std::barrier waits (CTA-wide / per 32 consecutive threads);
//   * dynamic shared memory is one 16-byte-aligned buffer per CTA;
//   * the few CUDA runtime calls the driver makes become no-ops or libc calls.
//
// What this checks: arithmetic, indexing, the shared-memory protocol between lanes (barrier placement included: a
// missing __syncwarp shows up as a data race / wrong sums here as it would on the GPU) and the driver's buffer wiring.
// What it cannot check: launch limits, alignment faults, occupancy.  tests/test_training_gpu.py covers the real thing.
#pragma once
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#define B200SPAI_HOST_EMU 1
#define __device__
#define __global__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __restrict__
#define __align__(n) __attribute__((aligned(n)))
#define __shared__ static  /* a static array of a kernel (template instance) is shared by the host threads of the CTA that
                              runs; CTAs run one after the other, so it behaves like per-CTA shared memory */

struct float2 { float x, y; };
struct alignas(16) float4 { float x, y, z, w; };
struct int4 { int x, y, z, w; };
static inline float2 make_float2(float x, float y) { return float2{x, y}; }
static inline float4 make_float4(float x, float y, float z, float w) { return float4{x, y, z, w}; }
struct emu_dim3 { unsigned x = 1, y = 1, z = 1; };
static thread_local emu_dim3 threadIdx, blockIdx;
static emu_dim3 blockDim, gridDim;

struct EmuCta {
  std::unique_ptr<std::barrier<>> cta;
  std::vector<std::unique_ptr<std::barrier<>>> warps;
  float* smem = nullptr;
};
static EmuCta* emu_cta = nullptr;
static inline void __syncthreads() { emu_cta->cta->arrive_and_wait(); }
static inline void __syncwarp() { emu_cta->warps[threadIdx.x >> 5]->arrive_and_wait(); }
template <typename T>
static inline T __ldg(const T* p) { return *p; }
static inline float __fadd_rn(float a, float b) { return a + b; }
static inline float __fmul_rn(float a, float b) { return a * b; }
#define B2_DYN_SMEM(name) float* name = emu_cta->smem

template <typename K, typename... A>
static void emu_launch(K kernel, long long grid, int block, size_t smem_bytes, A... args) {
  gridDim.x = (unsigned)grid;
  blockDim.x = (unsigned)block;
  for (long long b = 0; b < grid; ++b) {
    EmuCta cta;
    cta.cta = std::make_unique<std::barrier<>>(block);
    for (int w = 0; w < (block + 31) / 32; ++w) cta.warps.push_back(std::make_unique<std::barrier<>>(std::min(32, block - 32 * w)));
    void* raw = std::aligned_alloc(128, ((smem_bytes + 127) / 128 + 1) * 128);
    std::memset(raw, 0xff, ((smem_bytes + 127) / 128 + 1) * 128);  // NaN pattern: uninitialised shared memory must not go unnoticed
    cta.smem = static_cast<float*>(raw);
    emu_cta = &cta;
    std::vector<std::thread> th;
    th.reserve(block);
    for (int t = 0; t < block; ++t)
      th.emplace_back([=]() {
        threadIdx.x = (unsigned)t;
        blockIdx.x = (unsigned)b;
        kernel(args...);
      });
    for (auto& x : th) x.join();
    emu_cta = nullptr;
    std::free(raw);
  }
}
#define B2_LAUNCH(kernel, grid, block, smem, stream, ...) emu_launch(kernel, (long long)(grid), (int)(block), (size_t)(smem), __VA_ARGS__)

// ---- the slice of the CUDA runtime / library helpers the driver touches ------------------------------------------------
typedef void* cudaStream_t;
typedef int cudaError_t;
enum { cudaSuccess = 0 };
enum { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
template <typename K>
static inline cudaError_t cudaFuncSetAttribute(K, int, int) { return cudaSuccess; }
template <typename K>
static inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int* occ, K, int, size_t) { *occ = 2; return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { std::memset(p, v, n); return cudaSuccess; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline const char* cudaGetErrorString(cudaError_t) { return "emu"; }
enum { B200SPAI_ERR_CUDA = -1, B200SPAI_ERR_ARG = -2 };
static int emu_sm_count = 3;  // a small "device": several CTAs per kernel, several tiles per warp
namespace b200spai {
static inline int sm_count() { return emu_sm_count; }
static inline void count_launches(long long) {}
static inline void dev_free(void*) {}
static inline void set_error(const std::string& s) { std::fprintf(stderr, "b200spai emu: %s\n", s.c_str()); }
}  // namespace b200spai
#define B2_CUDA(call) do { if ((call) != cudaSuccess) return B200SPAI_ERR_CUDA; } while (0)
#define B2_TRY(expr) do { int _rc = (expr); if (_rc != 0) return _rc; } while (0)
#define B2_ARG(cond, msg) do { if (!(cond)) { ::b200spai::set_error(msg); return B200SPAI_ERR_ARG; } } while (0)
#define B2_LAUNCH_CHECK() do { } while (0)
