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

struct EmuWarpXch {
  alignas(16) uint64_t w[32][4];  // one exchange slot per lane (shuffles: w[lane][0]; mma: A, B fragments)
};
struct EmuCta {
  std::unique_ptr<std::barrier<>> cta;
  std::vector<std::unique_ptr<std::barrier<>>> warps;
  std::vector<EmuWarpXch> xch;
  float* smem = nullptr;
};
static EmuCta* emu_cta = nullptr;
static inline void __syncthreads() { emu_cta->cta->arrive_and_wait(); }
static inline void __syncwarp() { emu_cta->warps[threadIdx.x >> 5]->arrive_and_wait(); }
template <typename T>
static inline T __ldg(const T* p) { return *p; }
static inline uint32_t __float_as_uint(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float __uint_as_float(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// Warp shuffles: every lane publishes its value in the warp's exchange block, the warp meets, every lane reads its source
// lane, the warp meets again (the block is free for the next collective).  All 32 lanes must call (full mask).
template <typename T>
static inline T emu_shfl(T v, int src_lane) {
  static_assert(sizeof(T) <= 8, "shuffle of at most 8 bytes");
  EmuWarpXch& x = emu_cta->xch[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  std::memcpy(&x.w[lane][0], &v, sizeof(T));
  __syncwarp();
  T out;
  std::memcpy(&out, &x.w[src_lane & 31][0], sizeof(T));
  __syncwarp();
  return out;
}
template <typename T>
static inline T __shfl_xor_sync(unsigned, T v, int lane_mask, int width = 32) { (void)width; return emu_shfl(v, (int)(threadIdx.x & 31) ^ lane_mask); }
template <typename T>
static inline T __shfl_sync(unsigned, T v, int src, int width = 32) {
  const int lane = threadIdx.x & 31;
  return emu_shfl(v, (lane & ~(width - 1)) | (src & (width - 1)));
}

// mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 in software, with the fragment layout of the PTX ISA (lane = 4 g + t):
//   A (16 x 8, row):  a0 = (g, t)  a1 = (g + 8, t)  a2 = (g, t + 4)  a3 = (g + 8, t + 4)
//   B (8 x 8, col):   b0 = (k = t, n = g)  b1 = (k = t + 4, n = g)
//   C / D (16 x 8):   c0 = (g, 2t)  c1 = (g, 2t + 1)  c2 = (g + 8, 2t)  c3 = (g + 8, 2t + 1)
// Operands are taken as TF32 (low 13 mantissa bits ignored, as the hardware does); the 8 products of an output are summed
// in double and added to the fp32 accumulator with one rounding (the hardware's internal order is unspecified).
static inline void emu_mma_m16n8k8_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  EmuWarpXch& x = emu_cta->xch[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  uint32_t* mine = reinterpret_cast<uint32_t*>(&x.w[lane][0]);
  for (int i = 0; i < 4; ++i) mine[i] = a[i] & 0xffffe000u;
  mine[4] = b0 & 0xffffe000u;
  mine[5] = b1 & 0xffffe000u;
  __syncwarp();
  auto A = [&](int row, int k) {  // element (row, k) of the 16 x 8 A tile
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&x.w[4 * (row & 7) + (k & 3)][0]);
    return (double)__uint_as_float(w[(row >> 3) + 2 * (k >> 2)]);
  };
  auto B = [&](int k, int n) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&x.w[4 * n + (k & 3)][0]);
    return (double)__uint_as_float(w[4 + (k >> 2)]);
  };
  float out[4];
  for (int i = 0; i < 4; ++i) {
    const int row = g + 8 * (i >> 1), n = 2 * t + (i & 1);
    double sum = 0.0;
    for (int k = 0; k < 8; ++k) sum += A(row, k) * B(k, n);
    out[i] = (float)((double)c[i] + sum);
  }
  __syncwarp();
  for (int i = 0; i < 4; ++i) c[i] = out[i];
}
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
    cta.xch.resize((block + 31) / 32);
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
