// Backward pass of the GNN (training through the fused kernels; included by gnn.cu).
//
// What the reference gets from torch autograd for NodeEdgeProcessing (neural_cg/nn/gnns.py:77-97, MPLayer
// basic_layers.py:193-225, FeedForward :103-109) when workspace.training_step (workspace.py:96-112,178-180) calls
// loss.backward(): the gradient of every Linear / LayerNorm parameter given dL/dG.
//
// Design.  Same data layout and thread mapping as the forward FFMA kernels (thread = edge in target-sorted order, or
// thread = node; weights broadcast from shared memory).  The forward pass for training (gnn_train.cuh) keeps only the
// per-layer node features, edge features and aggregated messages; every MLP-internal value (LayerNorm statistics,
// pre-activations) is RECOMPUTED here from those, so nothing of size [E,48] or [E,16] per Linear ever exists.
//
//   * input gradients: g_in[k] = sum_j W[k][j] g[j] per thread (the transposed dense16, same broadcast LDS.128);
//   * weight gradients: dW[k][j] = sum over units of in[k] g[j] is a reduction over ALL edges.  A warp stages the
//     (in, g) vectors of its 32 units in a shared-memory tile and every lane forms 8 entries of the 16x16 outer-product
//     sum over the 32 units (lane -> row k = lane / 2, columns 8 (lane & 1) .. +7; 1 LDS.32 + 2 LDS.128 per 8 FMA),
//     adding them into the warp's private accumulator block in shared memory, which has the layout of the weight
//     blob.  At the end the CTA adds its warps' blocks in warp order and writes one partial per CTA; a second kernel
//     adds the partials in CTA order.  No atomics anywhere: the gradient is a deterministic function of the inputs
//     and the grid;
//   * node gradients: an edge writes its contribution to the gradient of x[dst] by position (a target's edges are
//     contiguous) and of x[src] by original edge id (the row-sorted view of the graph lists a source's edges); a
//     thread-per-node kernel adds them in ascending order.
//
// GELU is nn.GELU()'s x Phi(x) with derivative Phi(x) + x phi(x), both from the rational erfc the forward kernels use
// (gelu_both: |error| <= 1.5e-7, one rcp + one ex2 for the pair).  The values saved by the training forward come from the
// inference kernels (tensor-core 3xTF32), which agree with the FFMA arithmetic used here to ~1e-6 relative -- the
// measured gradient error against the reference is <= 1e-5 of max|grad| (bar 1e-3).
#pragma once

// The kernels of this file also compile as plain C++ (tests/host_emu: one host thread per CUDA thread, barriers for
// __syncwarp / __syncthreads) so that their arithmetic and indexing are checked against torch autograd on the CPU.
#ifndef B2_DYN_SMEM
#define B2_DYN_SMEM(name) extern __shared__ __align__(16) float name[]
#endif
#ifndef B2_LAUNCH
#define B2_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif

namespace b200spai {

constexpr int BW_WARPS = 4;
constexpr int BW_THREADS = 32 * BW_WARPS;
constexpr int SCR_LD = 20;                      // row stride (floats) of the staging tiles: 16-byte aligned rows,
                                                // conflict-free float4 stores (20 l mod 32 distinct per quarter warp)
constexpr int SCR_FLOATS = 2 * 32 * SCR_LD;     // `in` tile + `g` tile of one warp

// GELU and its derivative in one pass: gelu(x) = x Phi(x), gelu'(x) = Phi(x) + x phi(x).  Phi from the rational erfc of
// Abramowitz & Stegun 7.1.26 (the forward kernels' gelu_as, |error| <= 1.5e-7); its exp(-x^2 / 2) IS sqrt(2 pi) phi(x), so
// one rcp + one ex2 serve both values (erff + expf + erff would be ~90 instructions per element, this is ~22).
__device__ __forceinline__ void gelu_both(float x, float& act, float& grad) {
  const float y = fabsf(x) * 0.70710678118654752440f;
  float t, e;
#ifdef B200SPAI_HOST_EMU
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
  const float h = 0.5f * (p * t) * e;          // erfc(|x| / sqrt 2) / 2
  const float cdf = x >= 0.0f ? 1.0f - h : h;  // Phi(x)
  act = x * cdf;
  grad = fmaf(x * 0.39894228040143267794f, e, cdf);
}

// FeedForward forward that keeps the two pre-activations (what the backward needs besides the input).
template <int K>
__device__ __forceinline__ void mlp3_fwd_tape(const float* W, const float (&x)[K], float (&a1)[F], float (&a2)[F],
                                              float (&out)[F]) {
  float h[F];
  dense16<K>(W, W + K * F, x, a1);
#pragma unroll
  for (int j = 0; j < F; ++j) h[j] = gelu_as(a1[j]);
  const float* W2 = W + K * F + F;
  dense16<F>(W2, W2 + F * F, h, a2);
#pragma unroll
  for (int j = 0; j < F; ++j) h[j] = gelu_as(a2[j]);
  const float* W3 = W2 + F * F + F;
  dense16<F>(W3, W3 + F * F, h, out);
}

// gin[k] (+)= sum_j W[k][j] g[j]   (W [K][16] in shared memory, broadcast LDS.128; packed FMA: two partial sums per
// instruction, added at the end)
template <int K, bool ACC>
__device__ __forceinline__ void dense16_T(const float* W, const float (&g)[F], float (&gin)[K]) {
  float2 g2[8];
#pragma unroll
  for (int q = 0; q < 8; ++q) g2[q] = make_float2(g[2 * q], g[2 * q + 1]);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const float4* w4 = reinterpret_cast<const float4*>(W + k * F);
    float2 s = make_float2(ACC ? gin[k] : 0.0f, 0.0f);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 w = w4[q];
      s = ffma2(make_float2(w.x, w.y), g2[2 * q], s);
      s = ffma2(make_float2(w.z, w.w), g2[2 * q + 1], s);
    }
    gin[k] = s.x + s.y;
  }
}

// accW[k][j] += sum over the warp's 32 units of in[OFF + k] g[j]  (k, j < 16);  accB[j] += sum of g[j] (accB may be
// NULL).  `scr` is the warp's staging tile, accW / accB its private accumulators.  Must be called by all 32 lanes.
template <int K, int OFF>
__device__ __forceinline__ void wgrad16(float* scr, int lane, const float (&in)[K], const float (&g)[F], float* accW,
                                        float* accB) {
  float* s_in = scr;
  float* s_g = scr + 32 * SCR_LD;
  __syncwarp();  // the previous call's reads of the tile are complete
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    *reinterpret_cast<float4*>(s_in + lane * SCR_LD + 4 * q) =
        make_float4(in[OFF + 4 * q], in[OFF + 4 * q + 1], in[OFF + 4 * q + 2], in[OFF + 4 * q + 3]);
    *reinterpret_cast<float4*>(s_g + lane * SCR_LD + 4 * q) = make_float4(g[4 * q], g[4 * q + 1], g[4 * q + 2], g[4 * q + 3]);
  }
  __syncwarp();
  const int k = lane >> 1, j0 = (lane & 1) * 8;
  float2 a[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) a[i] = make_float2(0.0f, 0.0f);
#pragma unroll 8
  for (int l = 0; l < 32; ++l) {
    const float xi = s_in[l * SCR_LD + k];
    const float2 xx = make_float2(xi, xi);
    const float4 g0 = *reinterpret_cast<const float4*>(s_g + l * SCR_LD + j0);
    const float4 g1 = *reinterpret_cast<const float4*>(s_g + l * SCR_LD + j0 + 4);
    a[0] = ffma2(xx, make_float2(g0.x, g0.y), a[0]);
    a[1] = ffma2(xx, make_float2(g0.z, g0.w), a[1]);
    a[2] = ffma2(xx, make_float2(g1.x, g1.y), a[2]);
    a[3] = ffma2(xx, make_float2(g1.z, g1.w), a[3]);
  }
  // the lane's 8 accumulator entries are 32 contiguous, 16-byte aligned bytes: two 128-bit updates, conflict-free across
  // the warp (scalar updates at a stride of 8 floats were 8-way bank conflicts)
  float4* dst = reinterpret_cast<float4*>(accW + k * F + j0);
  float4 d0 = dst[0], d1 = dst[1];
  d0.x += a[0].x;
  d0.y += a[0].y;
  d0.z += a[1].x;
  d0.w += a[1].y;
  d1.x += a[2].x;
  d1.y += a[2].y;
  d1.z += a[3].x;
  d1.w += a[3].y;
  dst[0] = d0;
  dst[1] = d1;
  if (accB != nullptr && lane < F) {
    float s = 0.0f;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) s += s_g[l * SCR_LD + lane];
    accB[lane] += s;
  }
}

template <int K, int C>
struct WgradChunks {
  __device__ static __forceinline__ void run(float* scr, int lane, const float (&x)[K], const float (&g)[F], float* accW) {
    WgradChunks<K, C - 1>::run(scr, lane, x, g, accW);
    wgrad16<K, 16 * (C - 1)>(scr, lane, x, g, accW + 16 * (C - 1) * F, nullptr);
  }
};
template <int K>
struct WgradChunks<K, 0> {
  __device__ static __forceinline__ void run(float*, int, const float (&)[K], const float (&)[F], float*) {}
};

// Backward of one FeedForward (3 Linear, GELU between): input x (after the pre-norm, if any), the two pre-activations
// of the forward, gout = dL/d out.  Adds the weight gradients into `acc` (layout of the weight block: W1[K][16], b1,
// W2[16][16], b2, W3[16][16], b3) and returns gin = dL/dx.
template <int K, bool ACC = false>
__device__ __forceinline__ void mlp3_bwd(const float* W, float* acc, float* scr, int lane, const float (&x)[K],
                                         const float (&a1)[F], const float (&a2)[F], const float (&gout)[F],
                                         float (&gin)[K]) {
  const float* W2 = W + K * F + F;
  const float* W3 = W2 + F * F + F;
  float* A2 = acc + K * F + F;
  float* A3 = A2 + F * F + F;
  float h[F], d[F], g2[F], g1[F];
#pragma unroll
  for (int j = 0; j < F; ++j) gelu_both(a2[j], h[j], d[j]);
  wgrad16<F, 0>(scr, lane, h, gout, A3, A3 + F * F);
  dense16_T<F, false>(W3, gout, g2);
#pragma unroll
  for (int j = 0; j < F; ++j) g2[j] *= d[j];
#pragma unroll
  for (int j = 0; j < F; ++j) gelu_both(a1[j], h[j], d[j]);
  wgrad16<F, 0>(scr, lane, h, g2, A2, A2 + F * F);
  dense16_T<F, false>(W2, g2, g1);
#pragma unroll
  for (int j = 0; j < F; ++j) g1[j] *= d[j];
  WgradChunks<K, K / 16>::run(scr, lane, x, g1, acc);
  // bias of the first Linear: one more pass over the staged g tile of the last chunk call
  if (lane < F) {
    float s = 0.0f;
    const float* s_g = scr + 32 * SCR_LD;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) s += s_g[l * SCR_LD + lane];
    acc[K * F + lane] += s;
  }
  dense16_T<K, ACC>(W, g1, gin);   // ACC: added to what the caller already holds in gin
}

// ---- shared-memory frame of a backward kernel: [weights nW][BW_WARPS accumulators of nW][BW_WARPS staging tiles] ----
__device__ __forceinline__ void bwd_frame_init(float* smem, int nW, const float* w0, int n0, const float* w1, int n1) {
  for (int i = threadIdx.x; i < n0; i += BW_THREADS) smem[i] = w0[i];
  for (int i = threadIdx.x; i < n1; i += BW_THREADS) smem[n0 + i] = w1[i];
  for (int i = threadIdx.x; i < BW_WARPS * nW; i += BW_THREADS) smem[nW + i] = 0.0f;
  __syncthreads();
}
__device__ __forceinline__ void bwd_frame_flush(const float* smem, int nW, float* partial) {
  __syncthreads();
  for (int i = threadIdx.x; i < nW; i += BW_THREADS) {
    float s = 0.0f;
#pragma unroll
    for (int w = 0; w < BW_WARPS; ++w) s += smem[nW + w * nW + i];
    partial[(long long)blockIdx.x * nW + i] = s;
  }
}
static inline size_t bwd_frame_bytes(int nW) { return ((size_t)nW * (1 + BW_WARPS) + (size_t)BW_WARPS * SCR_FLOATS) * sizeof(float); }

// grad[dst(i)] = sum over CTAs (ascending) of partial[c][i]; i < len0 goes to dst0 + i, the rest to dst1 + (i - len0)
__global__ void __launch_bounds__(256) gnn_reduce_partials_kernel(const float* __restrict__ partial, int ncta, int nW,
                                                                  float* __restrict__ grad, long long dst0, int len0,
                                                                  long long dst1) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nW) return;
  float s = 0.0f;
  for (int c = 0; c < ncta; ++c) s += partial[(long long)c * nW + i];
  grad[i < len0 ? dst0 + i : dst1 + (i - len0)] = s;
}

// plain (coherent) loads: for buffers the same kernel also writes
__device__ __forceinline__ void load16p(const float* p, float* dst) {
  const float4* p4 = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 v = p4[q];
    dst[4 * q + 0] = v.x;
    dst[4 * q + 1] = v.y;
    dst[4 * q + 2] = v.z;
    dst[4 * q + 3] = v.w;
  }
}
__device__ __forceinline__ void store16(float* p, const float* v) {
  float4* p4 = reinterpret_cast<float4*>(p);
#pragma unroll
  for (int q = 0; q < 4; ++q) p4[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
}

// ---- decoder backward (gnns.py:88-95) ------------------------------------------------------------------------------
struct DecBwdParams {
  const int *src, *dst, *perm;
  long long E;
  const float* ef;   // edge features after the last MP layer (tiled layout)
  const float* x;    // node features after the last MP layer
  const float* w;    // decoder weights
  const float* gG;   // dL/dG [E][bb], caller's edge order
  int bb;
  float* ge;         // out: dL/d edge features [E][16] by position
  float* gxs;        // out: contribution to dL/dx[src] [E][16] by ORIGINAL edge id
  float* gxd;        // out: contribution to dL/dx[dst] [E][16] by position
  float* partial;
};

__global__ void __launch_bounds__(BW_THREADS) gnn_dec_bwd_kernel(const DecBwdParams P) {
  B2_DYN_SMEM(smem);
  bwd_frame_init(smem, MLP48, P.w, MLP48, nullptr, 0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* acc = smem + MLP48 + warp * MLP48;
  float* scr = smem + MLP48 * (1 + BW_WARPS) + warp * SCR_FLOATS;
  const long long ntiles = (P.E + 31) >> 5;
  for (long long tile = (long long)blockIdx.x * BW_WARPS + warp; tile < ntiles; tile += (long long)gridDim.x * BW_WARPS) {
    const long long p0 = tile * 32 + lane;
    const bool valid = p0 < P.E;
    const long long p = valid ? p0 : P.E - 1;
    float v[48], a1[F], a2[F], o[F], gout[F], gin[48];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 t = *ef_chunk(const_cast<float*>(P.ef), p, q);
      v[4 * q + 0] = t.x;
      v[4 * q + 1] = t.y;
      v[4 * q + 2] = t.z;
      v[4 * q + 3] = t.w;
    }
    load16(P.x + (long long)__ldg(P.src + p) * F, v + 16);
    load16(P.x + (long long)__ldg(P.dst + p) * F, v + 32);
    mlp3_fwd_tape<48>(smem, v, a1, a2, o);
    const long long eo = __ldg(P.perm + p);
#pragma unroll
    for (int j = 0; j < F; ++j) gout[j] = (valid && j < P.bb) ? __ldg(P.gG + eo * P.bb + j) : 0.0f;
    mlp3_bwd<48>(smem, acc, scr, lane, v, a1, a2, gout, gin);
    if (valid) {
      store16(P.ge + p * F, gin);
      store16(P.gxs + eo * F, gin + 16);
      store16(P.gxd + p * F, gin + 32);
    }
  }
  bwd_frame_flush(smem, MLP48, P.partial);
}

// ---- gradient of the node features: base + sum over the edges that read the node as target / as source -------------
__global__ void __launch_bounds__(256) gnn_node_accum_kernel(const int* __restrict__ segptr, const int* __restrict__ rsegptr,
                                                             const int* __restrict__ rperm, long long N,
                                                             const float* __restrict__ gxd, const float* __restrict__ gxs,
                                                             const float* base, float* out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float a[F];
  if (base) load16p(base + i * F, a);
  else {
#pragma unroll
    for (int k = 0; k < F; ++k) a[k] = 0.0f;
  }
  for (int p = segptr[i]; p < segptr[i + 1]; ++p) {
    float t[F];
    load16(gxd + (long long)p * F, t);
#pragma unroll
    for (int k = 0; k < F; ++k) a[k] += t[k];
  }
  for (int q = rsegptr[i]; q < rsegptr[i + 1]; ++q) {
    float t[F];
    load16(gxs + (long long)rperm[q] * F, t);
#pragma unroll
    for (int k = 0; k < F; ++k) a[k] += t[k];
  }
  store16(out + i * F, a);
}

// ---- aggregated messages of a layer (the node MLP's input), kept for the backward ------------------------------------
__global__ void __launch_bounds__(256) gnn_agg_kernel(const int* __restrict__ segptr, long long N,
                                                      const float* __restrict__ msg, float* __restrict__ agg) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N) return;
  float a[F];
#pragma unroll
  for (int k = 0; k < F; ++k) a[k] = 0.0f;
  for (int p = segptr[i]; p < segptr[i + 1]; ++p) {  // the order of gnn_node_kernel
    float t[F];
    load16(msg + (long long)p * F, t);
#pragma unroll
    for (int k = 0; k < F; ++k) a[k] += t[k];
  }
  store16(agg + i * F, a);
}

// LayerNorm (no affine: folded into the following Linear) forward in place; returns rstd
template <int K>
__device__ __forceinline__ float layernorm_fwd(float (&v)[K]) {
  float mean = 0.0f;
#pragma unroll
  for (int k = 0; k < K; ++k) mean += v[k];
  mean *= (1.0f / K);
  float var = 0.0f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const float c = v[k] - mean;
    var = fmaf(c, c, var);
  }
  const float rstd = 1.0f / sqrtf(var * (1.0f / K) + 1e-5f);
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = (v[k] - mean) * rstd;
  return rstd;
}
// g (gradient w.r.t. the normalised values y) -> gradient w.r.t. the LayerNorm input, in place
template <int K>
__device__ __forceinline__ void layernorm_bwd(const float (&y)[K], float rstd, float (&g)[K]) {
  float m1 = 0.0f, m2 = 0.0f;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    m1 += g[k];
    m2 = fmaf(g[k], y[k], m2);
  }
  m1 *= (1.0f / K);
  m2 *= (1.0f / K);
#pragma unroll
  for (int k = 0; k < K; ++k) g[k] = rstd * (g[k] - m1 - y[k] * m2);
}

// ---- message-passing layer, edge part (basic_layers.py:207-222) ----------------------------------------------------
struct MpBwdParams {
  const int *src, *dst, *perm;
  long long E;
  const float* xin;        // node features entering the layer
  const float* ef_in;      // edge features entering the layer (tiled); unused when FIRST
  const float* edge_attr;  // FIRST: raw edge inputs [E][edge_in], caller's order
  int edge_in;
  const float* w_layer;    // msg MLP | edge MLP
  const float* w_eenc;     // FIRST: edge encoder
  const float* gagg;       // dL/d aggregated messages [N][16]
  float* ge;               // in: dL/d edge features leaving the layer; out: entering it (FIRST: leaving the encoder)
  float* gxs;
  float* gxd;
  float* partial;
};

template <bool FIRST>
__global__ void __launch_bounds__(BW_THREADS) gnn_mp_edge_bwd_kernel(const MpBwdParams P) {
  constexpr int NW = 2 * MLP48 + (FIRST ? MLP16 : 0);
  B2_DYN_SMEM(smem);
  bwd_frame_init(smem, NW, P.w_layer, 2 * MLP48, P.w_eenc, FIRST ? MLP16 : 0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* acc = smem + NW + warp * NW;
  float* scr = smem + NW * (1 + BW_WARPS) + warp * SCR_FLOATS;
  const long long ntiles = (P.E + 31) >> 5;
  for (long long tile = (long long)blockIdx.x * BW_WARPS + warp; tile < ntiles; tile += (long long)gridDim.x * BW_WARPS) {
    const long long p0 = tile * 32 + lane;
    const bool valid = p0 < P.E;
    const long long p = valid ? p0 : P.E - 1;
    const int d = __ldg(P.dst + p), s = __ldg(P.src + p);
    const long long eo = __ldg(P.perm + p);
    float v[48];
    load16(P.xin + (long long)d * F, v);
    load16(P.xin + (long long)s * F, v + 16);
    if (FIRST) {
      float a[F], t1[F], t2[F], e0[F];
#pragma unroll
      for (int k = 0; k < F; ++k) a[k] = (k < P.edge_in) ? __ldg(P.edge_attr + eo * P.edge_in + k) : 0.0f;
      mlp3_fwd_tape<F>(smem + 2 * MLP48, a, t1, t2, e0);
#pragma unroll
      for (int k = 0; k < F; ++k) v[32 + k] = e0[k];
    } else {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 t = *ef_chunk(const_cast<float*>(P.ef_in), p, q);
        v[32 + 4 * q + 0] = t.x;
        v[32 + 4 * q + 1] = t.y;
        v[32 + 4 * q + 2] = t.z;
        v[32 + 4 * q + 3] = t.w;
      }
    }
    const float rstd = layernorm_fwd<48>(v);
    float gn[48], ge_out[F];
#pragma unroll
    for (int k = 0; k < 48; ++k) gn[k] = 0.0f;
    load16p(P.ge + p * F, ge_out);
#pragma unroll 1
    for (int m = 0; m < 2; ++m) {  // m = 0: message MLP (gradient arrives through the aggregate), 1: edge MLP
      float a1[F], a2[F], o[F], gout[F];
      mlp3_fwd_tape<48>(smem + m * MLP48, v, a1, a2, o);
      if (m == 0) load16(P.gagg + (long long)d * F, gout);
      else {
#pragma unroll
        for (int k = 0; k < F; ++k) gout[k] = ge_out[k];
      }
      if (!valid) {
#pragma unroll
        for (int k = 0; k < F; ++k) gout[k] = 0.0f;
      }
      mlp3_bwd<48, true>(smem + m * MLP48, acc + m * MLP48, scr, lane, v, a1, a2, gout, gn);
    }
    layernorm_bwd<48>(v, rstd, gn);
#pragma unroll
    for (int k = 0; k < F; ++k) ge_out[k] += gn[32 + k];  // residual: e_out = e_in + edge_mlp(...)
    if (valid) {
      store16(P.gxd + p * F, gn);
      store16(P.gxs + eo * F, gn + 16);
      store16(P.ge + p * F, ge_out);
    }
    if (FIRST) {  // edge encoder (gnns.py:84): weight gradients only, the raw edge inputs need none
      float a[F], t1[F], t2[F], e0[F], gin[F];
#pragma unroll
      for (int k = 0; k < F; ++k) a[k] = (k < P.edge_in) ? __ldg(P.edge_attr + eo * P.edge_in + k) : 0.0f;
      mlp3_fwd_tape<F>(smem + 2 * MLP48, a, t1, t2, e0);
      if (!valid) {
#pragma unroll
        for (int k = 0; k < F; ++k) ge_out[k] = 0.0f;
      }
      mlp3_bwd<F>(smem + 2 * MLP48, acc + 2 * MLP48, scr, lane, a, t1, t2, ge_out, gin);
    }
  }
  bwd_frame_flush(smem, NW, P.partial);
}

// ---- node MLPs: the node part of a layer (LayerNorm16 + MLP on the aggregate, basic_layers.py:203-206) when LN, the
// node encoder (gnns.py:83) otherwise -----------------------------------------------------------------------------------
struct NodeBwdParams {
  long long N;
  const float* in;   // LN: aggregated messages [N][16]; else raw node inputs [N][in_dim]
  int in_dim;
  const float* gx;   // dL/d output of the MLP (= dL/dx leaving the layer: the residual passes it through unchanged)
  const float* w;
  float* gin;        // LN: dL/d aggregate [N][16]; else unused
  float* partial;
};

template <bool LN>
__global__ void __launch_bounds__(BW_THREADS) gnn_node_bwd_kernel(const NodeBwdParams P) {
  B2_DYN_SMEM(smem);
  bwd_frame_init(smem, MLP16, P.w, MLP16, nullptr, 0);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* acc = smem + MLP16 + warp * MLP16;
  float* scr = smem + MLP16 * (1 + BW_WARPS) + warp * SCR_FLOATS;
  const long long ntiles = (P.N + 31) >> 5;
  for (long long tile = (long long)blockIdx.x * BW_WARPS + warp; tile < ntiles; tile += (long long)gridDim.x * BW_WARPS) {
    const long long i0 = tile * 32 + lane;
    const bool valid = i0 < P.N;
    const long long i = valid ? i0 : P.N - 1;
    float a[F], a1[F], a2[F], o[F], gout[F], g[F];
    float rstd = 1.0f;
    if (LN) {
      load16(P.in + i * F, a);
      rstd = layernorm_fwd<F>(a);
    } else {
#pragma unroll
      for (int k = 0; k < F; ++k) a[k] = (k < P.in_dim) ? __ldg(P.in + i * P.in_dim + k) : 0.0f;
    }
    mlp3_fwd_tape<F>(smem, a, a1, a2, o);
    load16(P.gx + i * F, gout);
    if (!valid) {
#pragma unroll
      for (int k = 0; k < F; ++k) gout[k] = 0.0f;
    }
    mlp3_bwd<F>(smem, acc, scr, lane, a, a1, a2, gout, g);
    if (LN) {
      layernorm_bwd<F>(a, rstd, g);
      if (valid) store16(P.gin + i * F, g);
    }
  }
  bwd_frame_flush(smem, MLP16, P.partial);
}

// ---- gradient of GraphSpmv w.r.t. the matrix values (basic_layers.py:112-142) -----------------------------------------
// y[row] += A_e x[col]            ->  dA_e[a][c] (+)= gy[row][a] x[col][c]
// transpose: y[col] += A_e^T x[row]  ->  dA_e[a][c] (+)= x[row][a] gy[col][c]
__global__ void __launch_bounds__(256) graph_outer_f32_kernel(const int* __restrict__ src, const int* __restrict__ dst,
                                                              const int* __restrict__ perm, long long E, int b,
                                                              const float* __restrict__ gy, const float* __restrict__ x,
                                                              int transpose, int accumulate, float* __restrict__ gA) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= E) return;
  const long long r = src[p], c = dst[p];
  float* out = gA + (long long)perm[p] * b * b;
  const float* u = transpose ? x + r * b : gy + r * b;   // indexed by the block row a
  const float* w = transpose ? gy + c * b : x + c * b;   // indexed by the block column c
  for (int a = 0; a < b; ++a)
    for (int j = 0; j < b; ++j) {
      const float t = u[a] * w[j];
      out[a * b + j] = accumulate ? out[a * b + j] + t : t;
    }
}

}  // namespace b200spai
