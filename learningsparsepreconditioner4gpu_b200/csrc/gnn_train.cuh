// Training entry points of the GNN (included at the end of gnn.cu): the forward pass that keeps what the
// backward needs, and the backward driver over the kernels of gnn_bwd.cuh.
//
//   b200spai_gnn_forward_train  = b200spai_gnn_forward (the same inference kernels, stage by stage) + after every stage a
//                                 copy of the node features, the edge features and the aggregated messages into the tape
//   b200spai_gnn_backward       = dL/dG [E,b*b]  ->  dL/d(weight blob), in the layout of the blob b200spai_gnn_create takes
//
// Stands in for what torch autograd records and replays for NodeEdgeProcessing.forward (neural_cg/nn/gnns.py:77-97) inside
// SimpleTrainingWorkspace.training_step (neural_cg/workspace.py:96-112,178-180).
#pragma once

struct b200spai_gnn_tape_s {
  long long N = 0, E = 0, Epad = 0;
  int L = 0;
  float* xs = nullptr;    // [L+1][N][16]   node features after the encoder and after every MP layer
  float* es = nullptr;    // [L][Epad][16]  edge features after MP layer 1..L (tiled layout of the forward kernels)
  float* aggs = nullptr;  // [L][N][16]     aggregated messages of every layer
  float *ge = nullptr, *gxs = nullptr, *gxd = nullptr, *gx = nullptr, *gagg = nullptr, *partial = nullptr;
  int max_ctas = 0;
  bool filled = false;
};

namespace {

using namespace b200spai;

constexpr int BWD_NW_MAX = 2 * MLP48 + MLP16;

void tape_free(b200spai_gnn_tape_s* t) {
  if (!t) return;
  dev_free(t->xs);
  dev_free(t->es);
  dev_free(t->aggs);
  dev_free(t->ge);
  dev_free(t->gxs);
  dev_free(t->gxd);
  dev_free(t->gx);
  dev_free(t->gagg);
  dev_free(t->partial);
  delete t;
}

template <typename Kern>
int bwd_grid(Kern kernel, size_t smem_bytes, long long units, int max_ctas, int* grid_out) {
  B2_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  int occ = 0;
  B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, BW_THREADS, smem_bytes));
  if (occ < 1) occ = 1;
  long long grid = ((units + 31) / 32 + BW_WARPS - 1) / BW_WARPS;
  if (grid > (long long)sm_count() * occ) grid = (long long)sm_count() * occ;
  if (grid > max_ctas) grid = max_ctas;
  if (grid < 1) grid = 1;
  *grid_out = (int)grid;
  return 0;
}

int reduce_partials(const float* partial, int ncta, int nW, float* grad, long long dst0, int len0, long long dst1,
                    cudaStream_t st) {
  B2_LAUNCH(gnn_reduce_partials_kernel, (nW + 255) / 256, 256, 0, st, partial, ncta, nW, grad, dst0, len0, dst1);
  B2_LAUNCH_CHECK();
  return 0;
}

int node_accum(b200spai_graph_t g, const float* gxd, const float* gxs, const float* base, float* out, cudaStream_t st) {
  B2_LAUNCH(gnn_node_accum_kernel, (unsigned)((g->N + 255) / 256), 256, 0, st, g->segptr, g->rsegptr, g->rperm, g->N, gxd, gxs, base, out);
  B2_LAUNCH_CHECK();
  return 0;
}

int gnn_backward_impl(b200spai_gnn_t m, b200spai_graph_t g, b200spai_gnn_tape_s* t, const float* x_dev,
                      const float* edge_attr_dev, const float* gG, float* grad, cudaStream_t st) {
  const long long N = g->N, E = g->E;
  const int L = m->L;
  B2_CUDA(cudaMemsetAsync(grad, 0, (size_t)m->blob_len * sizeof(float), st));
  if (N == 0 || E == 0) return 0;  // G is empty: the loss does not depend on the weights
  const long long off_eenc = MLP16, off_layers = 2 * MLP16, off_dec = 2 * MLP16 + (long long)L * LAYER_FLOATS;
  const size_t xsz = (size_t)N * F, esz = (size_t)t->Epad * F;
  int grid = 0;
  long long launches = 0;
  {  // decoder
    DecBwdParams P{};
    P.src = g->src;
    P.dst = g->dst;
    P.perm = g->perm;
    P.E = E;
    P.ef = t->es + (size_t)(L - 1) * esz;
    P.x = t->xs + (size_t)L * xsz;
    P.w = m->w_dec();
    P.gG = gG;
    P.bb = m->bb;
    P.ge = t->ge;
    P.gxs = t->gxs;
    P.gxd = t->gxd;
    P.partial = t->partial;
    const size_t bytes = bwd_frame_bytes(MLP48);
    B2_TRY(bwd_grid(gnn_dec_bwd_kernel, bytes, E, t->max_ctas, &grid));
    B2_LAUNCH(gnn_dec_bwd_kernel, grid, BW_THREADS, bytes, st, P);
    B2_LAUNCH_CHECK();
    B2_TRY(reduce_partials(t->partial, grid, MLP48, grad, off_dec, MLP48, 0, st));
    B2_TRY(node_accum(g, t->gxd, t->gxs, nullptr, t->gx, st));
    launches += 3;
  }
  for (int l = L; l >= 1; --l) {
    const long long off_l = off_layers + (long long)(l - 1) * LAYER_FLOATS;
    {  // node MLP of the layer: dL/dx_l -> dL/d aggregate
      NodeBwdParams P{};
      P.N = N;
      P.in = t->aggs + (size_t)(l - 1) * xsz;
      P.in_dim = F;
      P.gx = t->gx;
      P.w = m->w_layer(l - 1) + 2 * MLP48;
      P.gin = t->gagg;
      P.partial = t->partial;
      const size_t bytes = bwd_frame_bytes(MLP16);
      B2_TRY(bwd_grid(gnn_node_bwd_kernel<true>, bytes, N, t->max_ctas, &grid));
      B2_LAUNCH(gnn_node_bwd_kernel<true>, grid, BW_THREADS, bytes, st, P);
      B2_LAUNCH_CHECK();
      B2_TRY(reduce_partials(t->partial, grid, MLP16, grad, off_l + 2 * MLP48, MLP16, 0, st));
    }
    {  // message + edge MLPs
      MpBwdParams P{};
      P.src = g->src;
      P.dst = g->dst;
      P.perm = g->perm;
      P.E = E;
      P.xin = t->xs + (size_t)(l - 1) * xsz;
      P.ef_in = l >= 2 ? t->es + (size_t)(l - 2) * esz : nullptr;
      P.edge_attr = edge_attr_dev;
      P.edge_in = m->edge_in;
      P.w_layer = m->w_layer(l - 1);
      P.w_eenc = m->w_edge_enc();
      P.gagg = t->gagg;
      P.ge = t->ge;
      P.gxs = t->gxs;
      P.gxd = t->gxd;
      P.partial = t->partial;
      if (l == 1) {
        const int nW = 2 * MLP48 + MLP16;
        const size_t bytes = bwd_frame_bytes(nW);
        B2_TRY(bwd_grid(gnn_mp_edge_bwd_kernel<true>, bytes, E, t->max_ctas, &grid));
        B2_LAUNCH(gnn_mp_edge_bwd_kernel<true>, grid, BW_THREADS, bytes, st, P);
        B2_LAUNCH_CHECK();
        B2_TRY(reduce_partials(t->partial, grid, nW, grad, off_l, 2 * MLP48, off_eenc, st));
      } else {
        const int nW = 2 * MLP48;
        const size_t bytes = bwd_frame_bytes(nW);
        B2_TRY(bwd_grid(gnn_mp_edge_bwd_kernel<false>, bytes, E, t->max_ctas, &grid));
        B2_LAUNCH(gnn_mp_edge_bwd_kernel<false>, grid, BW_THREADS, bytes, st, P);
        B2_LAUNCH_CHECK();
        B2_TRY(reduce_partials(t->partial, grid, nW, grad, off_l, nW, 0, st));
      }
    }
    B2_TRY(node_accum(g, t->gxd, t->gxs, t->gx, t->gx, st));  // x_l = x_{l-1} + ...: the residual carries gx through
    launches += 5;
  }
  {  // node encoder
    NodeBwdParams P{};
    P.N = N;
    P.in = x_dev;
    P.in_dim = m->node_in;
    P.gx = t->gx;
    P.w = m->w_node_enc();
    P.gin = nullptr;
    P.partial = t->partial;
    const size_t bytes = bwd_frame_bytes(MLP16);
    B2_TRY(bwd_grid(gnn_node_bwd_kernel<false>, bytes, N, t->max_ctas, &grid));
    B2_LAUNCH(gnn_node_bwd_kernel<false>, grid, BW_THREADS, bytes, st, P);
    B2_LAUNCH_CHECK();
    B2_TRY(reduce_partials(t->partial, grid, MLP16, grad, 0, MLP16, 0, st));
    launches += 2;
  }
  m->launches += launches;
  count_launches(launches);
  return 0;
}

}  // namespace

#ifndef B200SPAI_HOST_EMU  // the C ABI needs the forward kernels and the CUDA runtime; the host build of the tests stops here
extern "C" {

int b200spai_gnn_tape_create(b200spai_gnn_t m, b200spai_graph_t g, b200spai_gnn_tape_t* out) {
  B2_ARG(m && g && out, "gnn_tape_create: NULL argument");
  auto* t = new b200spai_gnn_tape_s();
  t->N = g->N;
  t->E = g->E;
  t->Epad = ((g->E + 31) / 32) * 32;
  t->L = m->L;
  t->max_ctas = sm_count() * 4;
  const size_t xsz = (size_t)t->N * F, esz = (size_t)t->Epad * F;
  int rc = dev_alloc(&t->xs, (size_t)(t->L + 1) * xsz);
  if (!rc) rc = dev_alloc(&t->es, (size_t)t->L * esz);
  if (!rc) rc = dev_alloc(&t->aggs, (size_t)t->L * xsz);
  if (!rc) rc = dev_alloc(&t->ge, esz);
  if (!rc) rc = dev_alloc(&t->gxs, esz);
  if (!rc) rc = dev_alloc(&t->gxd, esz);
  if (!rc) rc = dev_alloc(&t->gx, xsz);
  if (!rc) rc = dev_alloc(&t->gagg, xsz);
  if (!rc) rc = dev_alloc(&t->partial, (size_t)t->max_ctas * BWD_NW_MAX);
  if (rc) {
    tape_free(t);
    return rc;
  }
  *out = t;
  return 0;
}

int b200spai_gnn_tape_destroy(b200spai_gnn_tape_t t) {
  tape_free(t);
  return 0;
}

int b200spai_gnn_forward_train(b200spai_gnn_t m, b200spai_graph_t g, b200spai_gnn_tape_t t, const float* x_dev,
                               const float* edge_attr_dev, float* g_out_dev, void* stream) {
  B2_ARG(m && g && t && x_dev && (edge_attr_dev || g->E == 0) && (g_out_dev || g->E == 0), "gnn_forward_train: NULL argument");
  B2_ARG(t->N == g->N && t->E == g->E && t->L == m->L, "gnn_forward_train: the tape was created for another graph / model");
  cudaStream_t st = as_stream(stream);
  t->filled = false;
  if (g->N == 0) {
    t->filled = true;
    return 0;
  }
  const size_t xsz = (size_t)g->N * F, esz = (size_t)t->Epad * F;
  for (int stage = 0; stage <= m->L + 1; ++stage) {
    B2_TRY(gnn_stage(m, g, stage, x_dev, edge_attr_dev, g_out_dev, st));
    if (stage > m->L) break;
    B2_CUDA(cudaMemcpyAsync(t->xs + (size_t)stage * xsz, (stage & 1) ? g->x1 : g->x0, xsz * sizeof(float),
                            cudaMemcpyDeviceToDevice, st));
    if (stage >= 1 && g->E > 0) {
      B2_CUDA(cudaMemcpyAsync(t->es + (size_t)(stage - 1) * esz, g->ef, esz * sizeof(float), cudaMemcpyDeviceToDevice, st));
      B2_LAUNCH(gnn_agg_kernel, (unsigned)((g->N + 255) / 256), 256, 0, st, g->segptr, g->N, g->msg, t->aggs + (size_t)(stage - 1) * xsz);
      B2_LAUNCH_CHECK();
      m->launches += 1;
      count_launches(1);
    }
  }
  t->filled = true;
  return 0;
}

int b200spai_gnn_backward(b200spai_gnn_t m, b200spai_graph_t g, b200spai_gnn_tape_t t, const float* x_dev,
                          const float* edge_attr_dev, const float* grad_g_dev, float* grad_blob_dev, void* stream) {
  B2_ARG(m && g && t && grad_blob_dev, "gnn_backward: NULL argument");
  B2_ARG(t->N == g->N && t->E == g->E && t->L == m->L, "gnn_backward: the tape was created for another graph / model");
  B2_ARG(t->filled, "gnn_backward: run b200spai_gnn_forward_train on this tape first");
  B2_ARG(g->E == 0 || (x_dev && edge_attr_dev && grad_g_dev), "gnn_backward: NULL argument");
  return gnn_backward_impl(m, g, t, x_dev, edge_attr_dev, grad_g_dev, grad_blob_dev, as_stream(stream));
}

int b200spai_gnn_set_weights_dev(b200spai_gnn_t m, const float* blob_dev, int64_t blob_len, void* stream) {
  B2_ARG(m && blob_dev, "gnn_set_weights_dev: NULL argument");
  B2_ARG(blob_len == m->blob_len, "gnn_set_weights_dev: weight blob has the wrong length");
  B2_CUDA(cudaMemcpyAsync(m->blob, blob_dev, (size_t)blob_len * sizeof(float), cudaMemcpyDeviceToDevice, as_stream(stream)));
  return 0;
}

int b200spai_graph_outer_f32(b200spai_graph_t g, int32_t block_size, const float* gy_dev, const float* x_dev,
                             int32_t transpose, int32_t accumulate, float* grad_values_dev, void* stream) {
  B2_ARG(g && gy_dev && x_dev && (grad_values_dev || g->E == 0), "graph_outer_f32: NULL argument");
  B2_ARG(block_size >= 1 && block_size <= 4, "graph_outer_f32: block_size must be in [1,4]");
  if (g->E == 0) return 0;
  B2_LAUNCH(graph_outer_f32_kernel, (unsigned)((g->E + 255) / 256), 256, 0, as_stream(stream), 
      g->src, g->dst, g->perm, g->E, block_size, gy_dev, x_dev, transpose, accumulate, grad_values_dev);
  B2_LAUNCH_CHECK();
  count_launches(1);
  return 0;
}

}  // extern "C"
#endif  // B200SPAI_HOST_EMU
