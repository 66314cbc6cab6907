// Device code of the GNN (included by gnn.cu after the constants and structs).
#pragma once

namespace b200spai {

// nn.GELU() default = exact erf form (neural_cg/nn/basic_layers.py:21-22)
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

// The same function, x * Phi(x) with Phi(x) = erfc(-x / sqrt 2) / 2, from the rational erfc approximation of
// Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7 absolute on erfc, i.e. at the fp32 rounding level of the erff form,
// and without its cancellation in the negative tail): 2 MUFU + 13 FP instructions instead of ~24 for erff.
__device__ __forceinline__ float gelu_as(float x) {
  const float y = fabsf(x) * 0.70710678118654752440f;
  float t, e;
#ifdef B200SPAI_HOST_EMU  // host build of the backward kernels (tests/host_emu): exact reciprocal / exp2
  t = 1.0f / fmaf(0.3275911f, y, 1.0f);
  e = exp2f(y * y * -1.4426950408889634f);
#else
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f, y, 1.0f)));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(y * y * -1.4426950408889634f));
#endif
  float p = fmaf(1.061405429f, t, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float h = 0.5f * (p * t) * e;  // erfc(|x|/sqrt 2) / 2
  return x * (x >= 0.0f ? 1.0f - h : h);
}
template <int G>
__device__ __forceinline__ float gelu(float x) { return G ? gelu_as(x) : gelu_erf(x); }

// Packed fp32 FMA (Blackwell FFMA2): two independent fma.rn.f32 per instruction, same bits as two FFMA.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
#ifdef B200SPAI_HOST_EMU  // host build of the backward kernels (tests/host_emu): the same two rounded FMAs
  return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y));
#else
  float2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)),
        "l"(reinterpret_cast<const unsigned long long&>(c)));
  return d;
#endif
}

// Packed fp32 multiply (FMUL2): two independent mul.rn.f32 per instruction, same bits as two FMUL.
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
#ifdef B200SPAI_HOST_EMU
  return make_float2(a.x * b.x, a.y * b.y);
#else
  float2 d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(reinterpret_cast<unsigned long long&>(d))
      : "l"(reinterpret_cast<const unsigned long long&>(a)), "l"(reinterpret_cast<const unsigned long long&>(b)));
  return d;
#endif
}
// gelu_as on a pair: the SAME operation sequence per element (every packed op rounds like its scalar form), so the
// result has the bits of (gelu_as(x.x), gelu_as(x.y)) at about two thirds of the issue slots.
__device__ __forceinline__ float2 gelu_as2(float2 x) {
  const float2 y = fmul2(make_float2(fabsf(x.x), fabsf(x.y)), make_float2(0.70710678118654752440f, 0.70710678118654752440f));
  const float2 d = ffma2(make_float2(0.3275911f, 0.3275911f), y, make_float2(1.0f, 1.0f));
  const float2 a = fmul2(fmul2(y, y), make_float2(-1.4426950408889634f, -1.4426950408889634f));
  float2 t, e;
#ifdef B200SPAI_HOST_EMU
  t = make_float2(1.0f / d.x, 1.0f / d.y);
  e = make_float2(exp2f(a.x), exp2f(a.y));
#else
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(d.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(d.y));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.x) : "f"(a.x));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e.y) : "f"(a.y));
#endif
  float2 p = ffma2(make_float2(1.061405429f, 1.061405429f), t, make_float2(-1.453152027f, -1.453152027f));
  p = ffma2(p, t, make_float2(1.421413741f, 1.421413741f));
  p = ffma2(p, t, make_float2(-0.284496736f, -0.284496736f));
  p = ffma2(p, t, make_float2(0.254829592f, 0.254829592f));
  const float2 h = fmul2(fmul2(make_float2(0.5f, 0.5f), fmul2(p, t)), e);
  const float2 c = make_float2(x.x >= 0.0f ? 1.0f - h.x : h.x, x.y >= 0.0f ? 1.0f - h.y : h.y);
  return fmul2(x, c);
}

// y[j] = bias[j] + sum_k x[k] * W[k][j];  W [K][16] and bias [16] in shared memory (16-byte aligned).
// Every lane reads the same addresses (broadcast LDS.128), 4 weights per load, 8 FFMA2 (= 16 FMA) per k.
template <int K>
__device__ __forceinline__ void dense16(const float* W, const float* bias, const float (&x)[K], float (&y)[F]) {
  const float4* b4 = reinterpret_cast<const float4*>(bias);
  float2 acc[8];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 v = b4[q];
    acc[2 * q] = make_float2(v.x, v.y);
    acc[2 * q + 1] = make_float2(v.z, v.w);
  }
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const float4* w4 = reinterpret_cast<const float4*>(W + k * F);
    const float2 xx = make_float2(x[k], x[k]);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 w = w4[q];
      acc[2 * q] = ffma2(xx, make_float2(w.x, w.y), acc[2 * q]);
      acc[2 * q + 1] = ffma2(xx, make_float2(w.z, w.w), acc[2 * q + 1]);
    }
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    y[2 * q] = acc[q].x;
    y[2 * q + 1] = acc[q].y;
  }
}

// FeedForward with 3 Linear layers and GELU in between (basic_layers.py:103-109, num_layers = 2).
// Weight block layout: W1[K][16], b1[16], W2[16][16], b2[16], W3[16][16], b3[16].
template <int K, int G = 0>
__device__ __forceinline__ void mlp3(const float* W, const float (&x)[K], float (&out)[F]) {
  float h1[F], h2[F];
  dense16<K>(W, W + K * F, x, h1);
#pragma unroll
  for (int j = 0; j < F; ++j) h1[j] = gelu<G>(h1[j]);
  const float* W2 = W + K * F + F;
  dense16<F>(W2, W2 + F * F, h1, h2);
#pragma unroll
  for (int j = 0; j < F; ++j) h2[j] = gelu<G>(h2[j]);
  const float* W3 = W2 + F * F + F;
  dense16<F>(W3, W3 + F * F, h2, out);
}

__device__ __forceinline__ void load16(const float* p, float* dst) {
  const float4* p4 = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 v = __ldg(p4 + q);
    dst[4 * q + 0] = v.x;
    dst[4 * q + 1] = v.y;
    dst[4 * q + 2] = v.z;
    dst[4 * q + 3] = v.w;
  }
}

// tile-transposed edge-feature address of (edge position p, 4-float chunk q)
__device__ __forceinline__ float4* ef_chunk(float* ef, long long p, int q) {
  return reinterpret_cast<float4*>(ef) + (((p >> 5) * 4 + q) * 32 + (p & 31));
}

// ---- node encoder ----------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256) gnn_node_enc_kernel(const float* __restrict__ x, int node_in, long long N,
                                                           const float* __restrict__ w, float* __restrict__ xout) {
  __shared__ __align__(16) float s_w[MLP16];
  for (int i = threadIdx.x; i < MLP16; i += blockDim.x) s_w[i] = w[i];
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < N; i += stride) {
    float a[F], o[F];
#pragma unroll
    for (int k = 0; k < F; ++k) a[k] = (k < node_in) ? __ldg(x + i * node_in + k) : 0.0f;
    mlp3<F, G>(s_w, a, o);
    float4* o4 = reinterpret_cast<float4*>(xout + i * F);
#pragma unroll
    for (int q = 0; q < 4; ++q) o4[q] = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
  }
}

// ---- message-passing layer = edge kernel + node kernel ------------------------------------------------------------
// Edge kernel: thread per edge (target-sorted position p), grid-stride, no barrier inside the loop.  Gathers x[dst],
// x[src], e; LayerNorm(48) statistics shared by both MLPs (their affines are folded into W1/b1 on the host); msg MLP ->
// msg[p] (row-major [E][16], 64 contiguous bytes per thread); edge MLP -> e += ... in place (it sees the OLD node
// features, basic_layers.py:208).  Layer 0 fuses the edge encoder (gnns.py:84).
// Node kernel: thread per target node: ordered sum of its segment's messages (ascending edge order = the CPU
// scatter_add_ order of the reference, no atomics), LayerNorm(16), node MLP, residual.
// The messages make one round trip through HBM (128 B per edge and layer, ~3 ms at 216^3): an earlier single-kernel
// version kept them in shared memory but paid for it with CTA-wide barriers and a serial per-block node phase that
// left the FMA pipe at 36 % (profiles/r01_gnn_mp_ncu_full.txt).
struct MpParams {
  const int* src;
  const int* dst;
  const int* segptr;
  const int* perm;
  long long E, N;
  float* ef;
  float* msg;
  const float* xin;
  float* xout;
  const float* w_layer;
  const float* w_eenc;
  const float* edge_attr;
  int edge_in;
};

// G: GELU evaluation (0 = erff, 1 = rational erfc).
template <bool FIRST, int G>
__global__ void __launch_bounds__(MP_THREADS, 2) gnn_edge_kernel(const MpParams P) {
  __shared__ __align__(16) float s_w[2 * MLP48 + (FIRST ? MLP16 : 0)];
  const int tid = threadIdx.x;
  for (int i = tid; i < 2 * MLP48; i += MP_THREADS) s_w[i] = P.w_layer[i];
  if (FIRST)
    for (int i = tid; i < MLP16; i += MP_THREADS) s_w[2 * MLP48 + i] = P.w_eenc[i];
  __syncthreads();
  const long long stride = (long long)gridDim.x * MP_THREADS;
  for (long long p = (long long)blockIdx.x * MP_THREADS + tid; p < P.E; p += stride) {
    float v[48];  // [x_i = x[dst], x_j = x[src], e]  (basic_layers.py:216-222)
    const int d = __ldg(P.dst + p), s = __ldg(P.src + p);
    load16(P.xin + (long long)d * F, v);
    load16(P.xin + (long long)s * F, v + 16);
    if (FIRST) {
      float a[F], e0[F];
      const long long eo = __ldg(P.perm + p);
#pragma unroll
      for (int k = 0; k < F; ++k) a[k] = (k < P.edge_in) ? __ldg(P.edge_attr + eo * P.edge_in + k) : 0.0f;
      mlp3<F, G>(s_w + 2 * MLP48, a, e0);
#pragma unroll
      for (int k = 0; k < F; ++k) v[32 + k] = e0[k];
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 t = *ef_chunk(P.ef, p, q);
        v[32 + 4 * q + 0] = t.x;
        v[32 + 4 * q + 1] = t.y;
        v[32 + 4 * q + 2] = t.z;
        v[32 + 4 * q + 3] = t.w;
      }
    }
    float ev[F];
#pragma unroll
    for (int k = 0; k < F; ++k) ev[k] = v[32 + k];
    float mean = 0.0f;
#pragma unroll
    for (int k = 0; k < 48; ++k) mean += v[k];
    mean *= (1.0f / 48.0f);
    float var = 0.0f;
#pragma unroll
    for (int k = 0; k < 48; ++k) {
      const float c = v[k] - mean;
      var = fmaf(c, c, var);
    }
    const float rstd = 1.0f / sqrtf(var * (1.0f / 48.0f) + 1e-5f);
#pragma unroll
    for (int k = 0; k < 48; ++k) v[k] = (v[k] - mean) * rstd;
    // msg MLP, then edge MLP, through one copy of the MLP code (the weight offset changes)
#pragma unroll 1
    for (int m = 0; m < 2; ++m) {
      float o[F];
      mlp3<48, G>(s_w + m * MLP48, v, o);
      if (m == 0) {
        float4* mp = reinterpret_cast<float4*>(P.msg + p * F);
#pragma unroll
        for (int q = 0; q < 4; ++q) mp[q] = make_float4(o[4 * q], o[4 * q + 1], o[4 * q + 2], o[4 * q + 3]);
      } else {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *ef_chunk(P.ef, p, q) = make_float4(ev[4 * q] + o[4 * q], ev[4 * q + 1] + o[4 * q + 1],
                                              ev[4 * q + 2] + o[4 * q + 2], ev[4 * q + 3] + o[4 * q + 3]);
      }
    }
  }
}

// x_out = x_in + node_mlp(LayerNorm16(sum of the messages of the node))   (basic_layers.py:203-206,224-225)
template <int G>
__global__ void __launch_bounds__(256) gnn_node_kernel(const MpParams P) {
  __shared__ __align__(16) float s_w[MLP16];
  for (int i = threadIdx.x; i < MLP16; i += blockDim.x) s_w[i] = P.w_layer[2 * MLP48 + i];
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < P.N; i += stride) {
    float a[F];
#pragma unroll
    for (int k = 0; k < F; ++k) a[k] = 0.0f;
    const int sb = __ldg(P.segptr + i), se = __ldg(P.segptr + i + 1);
    for (int p = sb; p < se; ++p) {  // ascending edge order, one rounded add per message and channel
      const float4* mp = reinterpret_cast<const float4*>(P.msg + (long long)p * F);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 t = mp[q];
        a[4 * q + 0] += t.x;
        a[4 * q + 1] += t.y;
        a[4 * q + 2] += t.z;
        a[4 * q + 3] += t.w;
      }
    }
    float mean = 0.0f;
#pragma unroll
    for (int k = 0; k < F; ++k) mean += a[k];
    mean *= (1.0f / 16.0f);
    float var = 0.0f;
#pragma unroll
    for (int k = 0; k < F; ++k) {
      const float c = a[k] - mean;
      var = fmaf(c, c, var);
    }
    const float rstd = 1.0f / sqrtf(var * (1.0f / 16.0f) + 1e-5f);
#pragma unroll
    for (int k = 0; k < F; ++k) a[k] = (a[k] - mean) * rstd;
    float o[F];
    mlp3<F, G>(s_w, a, o);
    const float4* xi = reinterpret_cast<const float4*>(P.xin + i * F);
    float4* xo = reinterpret_cast<float4*>(P.xout + i * F);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 t = __ldg(xi + q);
      xo[q] = make_float4(t.x + o[4 * q], t.y + o[4 * q + 1], t.z + o[4 * q + 2], t.w + o[4 * q + 3]);
    }
  }
}

// ---- decoder ---------------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256, 2) gnn_dec_kernel(const int* __restrict__ src, const int* __restrict__ dst,
                                                          const int* __restrict__ perm, float* ef,
                                                          const float* __restrict__ x, const float* __restrict__ w,
                                                          long long E, int bb, float* __restrict__ gout) {
  __shared__ __align__(16) float s_w[MLP48];
  for (int i = threadIdx.x; i < MLP48; i += blockDim.x) s_w[i] = w[i];
  __syncthreads();
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < E; p += stride) {
    float v[48], o[F];  // [edge, x[edge_index[0]], x[edge_index[1]]]  (gnns.py:88-90)
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 t = *ef_chunk(ef, p, q);
      v[4 * q + 0] = t.x;
      v[4 * q + 1] = t.y;
      v[4 * q + 2] = t.z;
      v[4 * q + 3] = t.w;
    }
    load16(x + (long long)__ldg(src + p) * F, v + 16);
    load16(x + (long long)__ldg(dst + p) * F, v + 32);
    mlp3<48, G>(s_w, v, o);
    const long long eo = __ldg(perm + p);
#pragma unroll
    for (int j = 0; j < F; ++j)
      if (j < bb) gout[eo * bb + j] = o[j];
  }
}

// ---- GraphSpmv (neural_cg/nn/basic_layers.py:112-142) ----------------------------------------------
// y[t] = mask[t] * sum over the edges aggregated at target t of  A_e x[other]   (TRANSPOSE: A_e^T x[other]).
// One thread per scalar row (t, i); edges of a target are visited in ascending original order, i.e. the
// order in which the reference's CPU scatter_add_ accumulates them.  fp32 like the reference operator.
template <bool TRANSPOSE>
__global__ void __launch_bounds__(256) graph_spmv_f32_kernel(const int* __restrict__ segptr, const int* __restrict__ perm,
                                                             const int* __restrict__ other, long long N, int b,
                                                             const float* __restrict__ vals, const float* __restrict__ x,
                                                             const float* __restrict__ mask, float* __restrict__ y) {
  const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (row >= N * b) return;
  const long long t = row / b;
  const int i = (int)(row - t * b);
  float acc = 0.0f;
  for (int p = segptr[t]; p < segptr[t + 1]; ++p) {
    const float* blk = vals + (long long)perm[p] * b * b;
    const float* xo = x + (long long)other[p] * b;
    float r = 0.0f;
    for (int j = 0; j < b; ++j) r = __fadd_rn(r, __fmul_rn(TRANSPOSE ? blk[j * b + i] : blk[i * b + j], xo[j]));
    acc = __fadd_rn(acc, r);
  }
  y[row] = mask ? acc * mask[row] : acc;
}

}  // namespace b200spai
