// Per-edge MLPs of the GNN on the tensor cores (included by gnn.cu after gnn_kernels.cuh).
//
// Why: the thread-per-edge FFMA kernels feed a 48->16->16->16 GEMM through shared memory one edge per thread: 4 broadcast
// LDS.128 of weights per 8 FFMA2, l1tex shared-memory wavefronts at 83 % of peak, FMA pipe 55 %
// (profiles/r01_gnn_edge_ncu_full.txt) -- contraction-bound, which is the case the north star reserves tensor cores for.
//
// How: a warp owns a tile of 32 edges = two m16 row tiles of `mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32`.
// Every product is computed as 3 x TF32 with fp32 accumulation (error-compensated split):
//     a = a_hi + a_lo,  w = w_hi + w_lo  (hi = cvt.rna.tf32, lo = cvt.rna.tf32 of the exact remainder)
//     acc += a_lo * w_hi;  acc += a_hi * w_lo;  acc += a_hi * w_hi          (a_lo * w_lo ~ 2^-22 |a w| is dropped)
// which keeps |dG| / max|G| at the 1e-6 level of the FFMA kernels (the parity bar is 1e-5; single-pass TF32 misses it).
//
// Register-only layer chaining: the K order of every Linear is free (weight rows are permuted when the per-lane
// B-fragment tables are built), so the C fragment of one layer IS the A fragment of the next -- no shuffles, no shared
// memory round trip between layers:
//     lane = 4 g + t.   C: c0,c1 = row g, cols 2t,2t+1;  c2,c3 = row g+8.   A(chunk j): a0 = (g, col t), a1 = (g+8, col t),
//     a2 = (g, col t+4), a3 = (g+8, col t+4)   =>   chunk j := n-tile j,  col t := hidden 8j+2t,  col t+4 := hidden 8j+2t+1.
// Inputs: lane t loads features 4t..4t+3 of a 16-feature row as one float4 (4 lanes = one 64-byte row, 8 rows per
// instruction); block B, chunk c = 2B+j: col t := feature 16B+4t+2j, col t+4 := 16B+4t+2j+1.
// Outputs of the LAST Linear are permuted the other way (logical column 8nt+2t+i := feature 4t+2nt+i), so a lane ends
// up with features 4t..4t+3 of rows g and g+8: one float4 store per row, and the residual e += ... lines up with the
// float4 of e the lane loaded.
//
// Weights: per-lane B fragments {w_hi(b0), w_hi(b1), w_lo(b0), w_lo(b1)} as float4 [entry][lane] in shared memory
// (conflict-free LDS.128, one load feeds 3 mma x 2 row tiles), built once per CTA from the fp32 blob.
#pragma once

namespace b200spai {

constexpr int MMA_NE48 = 2 * 6 + 8;  // table entries of a 48->16->16->16 MLP (chunks x n-tiles per layer)
constexpr int MMA_NE16 = 2 * 2 + 8;  // 16->16->16->16

__device__ __forceinline__ uint32_t tf32_rna(float x) {
#ifdef B200SPAI_HOST_EMU  // host build (tests/host_emu): round to nearest, ties away, on the bit pattern (finite inputs)
  return (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
#else
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
#endif
}

// Activations: cvt.rna.tf32.f32 compiles to 4 instructions on sm_100a (FSETP |x| < inf, IADD 0x1000, SEL, LOP3); the
// activations are finite, so round-to-nearest (ties away) is done by hand in 2: add half a TF32 ulp to the bit pattern,
// clear the 13 low bits.  The exact remainder a - hi (one FADD) is rounded the same way.  volatile: the splits of the
// layer-1 inputs must not be hoisted out of the msg/edge loop (they would occupy 96 registers).
__device__ __forceinline__ void tf32_split(float a, uint32_t& hi, uint32_t& lo) {
  uint32_t h, l;
#ifdef B200SPAI_HOST_EMU
  h = __float_as_uint(a) + 0x1000u;
#else
  asm volatile("add.u32 %0, %1, 0x1000;" : "=r"(h) : "r"(__float_as_uint(a)));
#endif
  h &= 0xffffe000u;
  l = (__float_as_uint(a - __uint_as_float(h)) + 0x1000u) & 0xffffe000u;
  hi = h;
  lo = l;
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
#ifdef B200SPAI_HOST_EMU  // warp-collective software mma with the PTX fragment layout (tests/host_emu/cuda_emu.h)
  emu_mma_m16n8k8_tf32(c, a, b0, b1);
#else
  asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
#endif
}

// Shared-memory image of one 3-Linear MLP: B fragments + C-fragment-ordered biases.
template <int KC1>
struct MmaMlpTab {
  float4 w[(2 * KC1 + 8) * 32];
  float2 bias[3 * 2 * 32];
};

// Build the table from the blob block [W1[K][16], b1, W2[16][16], b2, W3[16][16], b3]  (K = 8 KC1); all threads of the CTA.
template <int KC1>
__device__ __forceinline__ void mma_build_tab(const float* __restrict__ W, MmaMlpTab<KC1>* T, int tid, int nthreads) {
  constexpr int K = 8 * KC1;
  const float* Wl[3] = {W, W + K * F + F, W + K * F + F + F * F + F};
  const float* bl[3] = {W + K * F, W + K * F + F + F * F, W + K * F + F + 2 * (F * F) + F};
  constexpr int NE = 2 * KC1 + 8;
  for (int e = tid; e < NE * 32; e += nthreads) {
    const int idx = e >> 5, lane = e & 31, g = lane >> 2, t = lane & 3;
    int layer, c, nt;
    if (idx < 2 * KC1) { layer = 0; c = idx >> 1; nt = idx & 1; }
    else if (idx < 2 * KC1 + 4) { layer = 1; c = (idx - 2 * KC1) >> 1; nt = idx & 1; }
    else { layer = 2; c = (idx - 2 * KC1 - 4) >> 1; nt = idx & 1; }
    const int k0 = layer == 0 ? 16 * (c >> 1) + 4 * t + 2 * (c & 1) : 8 * c + 2 * t;  // col t; col t+4 is k0 + 1
    const int n = layer == 2 ? 4 * (g >> 1) + 2 * nt + (g & 1) : 8 * nt + g;
    const float w0 = __ldg(Wl[layer] + k0 * F + n), w1 = __ldg(Wl[layer] + (k0 + 1) * F + n);
    const float h0 = __uint_as_float(tf32_rna(w0)), h1 = __uint_as_float(tf32_rna(w1));
    T->w[e] = make_float4(h0, h1, __uint_as_float(tf32_rna(w0 - h0)), __uint_as_float(tf32_rna(w1 - h1)));
  }
  for (int e = tid; e < 3 * 2 * 32; e += nthreads) {
    const int layer = e >> 6, nt = (e >> 5) & 1, t = e & 3;
    const int n0 = layer == 2 ? 4 * t + 2 * nt : 8 * nt + 2 * t;
    T->bias[e] = make_float2(__ldg(bl[layer] + n0), __ldg(bl[layer] + n0 + 1));
  }
}

// One Linear on fragments: acc[mt][nt] (C layout, preset to the bias) += A[mt][c] * B[c][nt], 3xTF32.
template <int KC, int MT>
__device__ __forceinline__ void mma_linear(const float4* __restrict__ tw, int lane, const float (&a)[MT][KC][4],
                                           float (&acc)[MT][2][4]) {
#pragma unroll
  for (int c = 0; c < KC; ++c) {
    uint32_t hi[MT][4], lo[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        tf32_split(a[mt][c][i], hi[mt][i], lo[mt][i]);
      }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const float4 w = tw[(c * 2 + nt) * 32 + lane];
      const uint32_t wh0 = __float_as_uint(w.x), wh1 = __float_as_uint(w.y), wl0 = __float_as_uint(w.z), wl1 = __float_as_uint(w.w);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_tf32(acc[mt][nt], lo[mt], wh0, wh1);
        mma_tf32(acc[mt][nt], hi[mt], wl0, wl1);
        mma_tf32(acc[mt][nt], hi[mt], wh0, wh1);
      }
    }
  }
}

template <int MT>
__device__ __forceinline__ void mma_bias(const float2* __restrict__ tb, int layer, int lane, float (&acc)[MT][2][4]) {
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    const float2 b = tb[(layer * 2 + nt) * 32 + lane];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
      acc[mt][nt][0] = b.x;
      acc[mt][nt][1] = b.y;
      acc[mt][nt][2] = b.x;
      acc[mt][nt][3] = b.y;
    }
  }
}

// GELU on a C fragment, re-labelled as the A fragment of the next Linear (chunk j = n-tile j).
template <int MT, int G>
__device__ __forceinline__ void mma_act(const float (&acc)[MT][2][4], float (&a)[MT][2][4]) {
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      if (G) {  // rational-erfc GELU on pairs (packed fp32 arithmetic, the bits of the scalar form)
        const float2 u = gelu_as2(make_float2(acc[mt][j][0], acc[mt][j][2]));
        const float2 v = gelu_as2(make_float2(acc[mt][j][1], acc[mt][j][3]));
        a[mt][j][0] = u.x;
        a[mt][j][1] = u.y;
        a[mt][j][2] = v.x;
        a[mt][j][3] = v.y;
      } else {
        a[mt][j][0] = gelu<G>(acc[mt][j][0]);
        a[mt][j][1] = gelu<G>(acc[mt][j][2]);
        a[mt][j][2] = gelu<G>(acc[mt][j][1]);
        a[mt][j][3] = gelu<G>(acc[mt][j][3]);
      }
    }
}

// FeedForward (3 Linear, GELU between) on fragments.  out[mt][h] = features 4t..4t+3 of row g + 8 h of row tile mt.
template <int KC1, int MT, int G>
__device__ __forceinline__ void mlp3_mma(const MmaMlpTab<KC1>* T, int lane, const float (&a)[MT][KC1][4], float4 (&out)[MT][2]) {
  float acc[MT][2][4], h[MT][2][4];
  mma_bias<MT>(T->bias, 0, lane, acc);
  mma_linear<KC1, MT>(T->w, lane, a, acc);
  mma_act<MT, G>(acc, h);
  mma_bias<MT>(T->bias, 1, lane, acc);
  mma_linear<2, MT>(T->w + 2 * KC1 * 32, lane, h, acc);
  mma_act<MT, G>(acc, h);
  mma_bias<MT>(T->bias, 2, lane, acc);
  mma_linear<2, MT>(T->w + (2 * KC1 + 4) * 32, lane, h, acc);
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    out[mt][0] = make_float4(acc[mt][0][0], acc[mt][0][1], acc[mt][1][0], acc[mt][1][1]);
    out[mt][1] = make_float4(acc[mt][0][2], acc[mt][0][3], acc[mt][1][2], acc[mt][1][3]);
  }
}

// The first Linear of TWO FeedForwards that read the same input (the message and the edge MLP of a layer share the
// LayerNorm'ed [x_i, x_j, e]): every input chunk is split into its TF32 halves ONCE and multiplied into both weight
// tables.  Per output the chunks and the three partial products arrive in the order of mma_linear: the same bits.
template <int KC, int MT>
__device__ __forceinline__ void mma_linear_pair(const float4* __restrict__ twA, const float4* __restrict__ twB, int lane,
                                                const float (&a)[MT][KC][4], float (&accA)[MT][2][4], float (&accB)[MT][2][4]) {
#pragma unroll
  for (int c = 0; c < KC; ++c) {
    uint32_t hi[MT][4], lo[MT][4];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
      for (int i = 0; i < 4; ++i) tf32_split(a[mt][c][i], hi[mt][i], lo[mt][i]);
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const float4 wa = twA[(c * 2 + nt) * 32 + lane], wb = twB[(c * 2 + nt) * 32 + lane];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        mma_tf32(accA[mt][nt], lo[mt], __float_as_uint(wa.x), __float_as_uint(wa.y));
        mma_tf32(accB[mt][nt], lo[mt], __float_as_uint(wb.x), __float_as_uint(wb.y));
        mma_tf32(accA[mt][nt], hi[mt], __float_as_uint(wa.z), __float_as_uint(wa.w));
        mma_tf32(accB[mt][nt], hi[mt], __float_as_uint(wb.z), __float_as_uint(wb.w));
        mma_tf32(accA[mt][nt], hi[mt], __float_as_uint(wa.x), __float_as_uint(wa.y));
        mma_tf32(accB[mt][nt], hi[mt], __float_as_uint(wb.x), __float_as_uint(wb.y));
      }
    }
  }
}

// Layers 2 and 3 of a FeedForward whose first Linear is already in `acc` (bias included).
template <int KC1, int MT, int G>
__device__ __forceinline__ void mlp3_mma_tail(const MmaMlpTab<KC1>* T, int lane, float (&acc)[MT][2][4], float4 (&out)[MT][2]) {
  float h[MT][2][4];
  mma_act<MT, G>(acc, h);
  mma_bias<MT>(T->bias, 1, lane, acc);
  mma_linear<2, MT>(T->w + 2 * KC1 * 32, lane, h, acc);
  mma_act<MT, G>(acc, h);
  mma_bias<MT>(T->bias, 2, lane, acc);
  mma_linear<2, MT>(T->w + (2 * KC1 + 4) * 32, lane, h, acc);
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) {
    out[mt][0] = make_float4(acc[mt][0][0], acc[mt][0][1], acc[mt][1][0], acc[mt][1][1]);
    out[mt][1] = make_float4(acc[mt][0][2], acc[mt][0][3], acc[mt][1][2], acc[mt][1][3]);
  }
}

// float4 of features 4t..4t+3 of the row held in half h -> the two input chunks 2B, 2B+1 of block B
template <int KC, int MT>
__device__ __forceinline__ void mma_put(float (&a)[MT][KC][4], int mt, int B, int h, const float4 f) {
  a[mt][2 * B][h] = f.x;
  a[mt][2 * B][h + 2] = f.y;
  a[mt][2 * B + 1][h] = f.z;
  a[mt][2 * B + 1][h + 2] = f.w;
}

// LayerNorm statistics of the rows a lane shares with the 3 other lanes of its quad (each lane holds KC*2 of the 8*KC
// features of row (mt, h)); normalises in place.
template <int KC, int MT>
__device__ __forceinline__ void mma_layernorm(float (&a)[MT][KC][4]) {
  constexpr float inv = 1.0f / (8 * KC);
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float s = 0.0f;
#pragma unroll
      for (int c = 0; c < KC; ++c) s += a[mt][c][h] + a[mt][c][h + 2];
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      const float mean = s * inv;
      float var = 0.0f;
#pragma unroll
      for (int c = 0; c < KC; ++c) {
        const float d0 = a[mt][c][h] - mean, d1 = a[mt][c][h + 2] - mean;
        var = fmaf(d0, d0, var);
        var = fmaf(d1, d1, var);
      }
      var += __shfl_xor_sync(0xffffffffu, var, 1);
      var += __shfl_xor_sync(0xffffffffu, var, 2);
      const float rstd = 1.0f / sqrtf(var * inv + 1e-5f);
#pragma unroll
      for (int c = 0; c < KC; ++c) {
        a[mt][c][h] = (a[mt][c][h] - mean) * rstd;
        a[mt][c][h + 2] = (a[mt][c][h + 2] - mean) * rstd;
      }
    }
}

constexpr int MMA_THREADS = 256;
constexpr int MMA_MT = 2;  // row tiles per warp: 32 edges

// ---- message-passing edge kernel on the tensor cores (same contract as gnn_edge_kernel) --------------------------
template <bool FIRST, int G>
__global__ void __launch_bounds__(MMA_THREADS, 2) gnn_edge_mma_kernel(const MpParams P) {
  __shared__ __align__(16) MmaMlpTab<6> s_msg, s_edge;
  __shared__ __align__(16) MmaMlpTab<2> s_enc;  // edge encoder (layer 0 only)
  const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, t = lane & 3;
  mma_build_tab<6>(P.w_layer, &s_msg, tid, MMA_THREADS);
  mma_build_tab<6>(P.w_layer + MLP48, &s_edge, tid, MMA_THREADS);
  if (FIRST) mma_build_tab<2>(P.w_eenc, &s_enc, tid, MMA_THREADS);
  __syncthreads();
  const long long ntiles = (P.E + 31) >> 5;
  const long long wstride = (long long)gridDim.x * (MMA_THREADS / 32);
  for (long long tile = (long long)blockIdx.x * (MMA_THREADS / 32) + (tid >> 5); tile < ntiles; tile += wstride) {
    const long long p0 = tile << 5;
    float v[MMA_MT][6][4];     // [x_i = x[dst], x_j = x[src], e]  (basic_layers.py:216-222) in A-fragment order
    float4 ev[MMA_MT][2];      // e of the lane's rows (residual)
    if (FIRST) {               // edge encoder e = MLP(edge_attr) on the fly (gnns.py:84)
      float a0[MMA_MT][2][4];
#pragma unroll
      for (int mt = 0; mt < MMA_MT; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const long long p = p0 + 16 * mt + 8 * h + g;
          const long long eo = p < P.E ? __ldg(P.perm + p) : 0;
          float f[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) f[i] = (p < P.E && 4 * t + i < P.edge_in) ? __ldg(P.edge_attr + eo * P.edge_in + 4 * t + i) : 0.0f;
          mma_put<2, MMA_MT>(a0, mt, 0, h, make_float4(f[0], f[1], f[2], f[3]));
        }
      mlp3_mma<2, MMA_MT, G>(&s_enc, lane, a0, ev);
    }
#pragma unroll
    for (int mt = 0; mt < MMA_MT; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const long long p = p0 + 16 * mt + 8 * h + g;
        const bool in = p < P.E;
        const int d = in ? __ldg(P.dst + p) : 0, s = in ? __ldg(P.src + p) : 0;
        mma_put<6, MMA_MT>(v, mt, 0, h, __ldg(reinterpret_cast<const float4*>(P.xin + (long long)d * F) + t));
        mma_put<6, MMA_MT>(v, mt, 1, h, __ldg(reinterpret_cast<const float4*>(P.xin + (long long)s * F) + t));
        // ef is padded to a whole tile; later layers re-read e for the residual (L1/L2 hit) instead of holding it
        mma_put<6, MMA_MT>(v, mt, 2, h, FIRST ? ev[mt][h] : *ef_chunk(P.ef, p, t));
      }
    mma_layernorm<6, MMA_MT>(v);
    // first Linear of the message and the edge MLP together (one TF32 split of the shared input), then the two tails
    float accM[MMA_MT][2][4], accE[MMA_MT][2][4];
    mma_bias<MMA_MT>(s_msg.bias, 0, lane, accM);
    mma_bias<MMA_MT>(s_edge.bias, 0, lane, accE);
    mma_linear_pair<6, MMA_MT>(s_msg.w, s_edge.w, lane, v, accM, accE);
    {
      float4 o[MMA_MT][2];
      mlp3_mma_tail<6, MMA_MT, G>(&s_msg, lane, accM, o);
#pragma unroll
      for (int mt = 0; mt < MMA_MT; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const long long p = p0 + 16 * mt + 8 * h + g;
          reinterpret_cast<float4*>(P.msg + p * F)[t] = o[mt][h];  // msg is padded to a whole tile
        }
    }
    {
      float4 o[MMA_MT][2];
      mlp3_mma_tail<6, MMA_MT, G>(&s_edge, lane, accE, o);
#pragma unroll
      for (int mt = 0; mt < MMA_MT; ++mt)
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const long long p = p0 + 16 * mt + 8 * h + g;
          const float4 e = FIRST ? ev[mt][h] : *ef_chunk(P.ef, p, t), r = o[mt][h];
          *ef_chunk(P.ef, p, t) = make_float4(e.x + r.x, e.y + r.y, e.z + r.z, e.w + r.w);
        }
    }
  }
}

// ---- decoder on the tensor cores (same contract as gnn_dec_kernel) ------------------------------------------------
template <int G>
__global__ void __launch_bounds__(MMA_THREADS, 2) gnn_dec_mma_kernel(const int* __restrict__ src, const int* __restrict__ dst,
                                                                     const int* __restrict__ perm, float* ef,
                                                                     const float* __restrict__ x, const float* __restrict__ w,
                                                                     long long E, int bb, float* __restrict__ gout) {
  __shared__ __align__(16) MmaMlpTab<6> s_dec;
  const int tid = threadIdx.x, lane = tid & 31, g = lane >> 2, t = lane & 3;
  mma_build_tab<6>(w, &s_dec, tid, MMA_THREADS);
  __syncthreads();
  const long long ntiles = (E + 31) >> 5;
  const long long wstride = (long long)gridDim.x * (MMA_THREADS / 32);
  for (long long tile = (long long)blockIdx.x * (MMA_THREADS / 32) + (tid >> 5); tile < ntiles; tile += wstride) {
    const long long p0 = tile << 5;
    float v[MMA_MT][6][4];  // [edge, x[edge_index[0]], x[edge_index[1]]]  (gnns.py:88-90)
#pragma unroll
    for (int mt = 0; mt < MMA_MT; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const long long p = p0 + 16 * mt + 8 * h + g;
        const bool in = p < E;
        const int d = in ? __ldg(dst + p) : 0, s = in ? __ldg(src + p) : 0;
        mma_put<6, MMA_MT>(v, mt, 0, h, *ef_chunk(ef, p, t));
        mma_put<6, MMA_MT>(v, mt, 1, h, __ldg(reinterpret_cast<const float4*>(x + (long long)s * F) + t));
        mma_put<6, MMA_MT>(v, mt, 2, h, __ldg(reinterpret_cast<const float4*>(x + (long long)d * F) + t));
      }
    float4 o[MMA_MT][2];
    mlp3_mma<6, MMA_MT, G>(&s_dec, lane, v, o);
#pragma unroll
    for (int mt = 0; mt < MMA_MT; ++mt)
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const long long p = p0 + 16 * mt + 8 * h + g;
        if (p < E && 4 * t < bb) {
          const long long eo = __ldg(perm + p);
          const float f[4] = {o[mt][h].x, o[mt][h].y, o[mt][h].z, o[mt][h].w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (4 * t + i < bb) gout[eo * bb + 4 * t + i] = f[i];
        }
      }
  }
}

}  // namespace b200spai
