// Constants and handle structs of the GNN (shared by gnn.cu and the host build of the backward kernels that
// tests/host_emu compiles with g++ to check their arithmetic on the CPU).
#pragma once

namespace b200spai {

constexpr int F = 16;                // hidden width (config/gnn.yaml gnn_features)
constexpr int MP_THREADS = 256;
constexpr int MLP16 = 3 * (F * F + F);        // 816 floats: three 16x16 Linear (+bias)
constexpr int MLP48 = 48 * F + F + 2 * (F * F + F);  // 1328 floats: 48->16, 16->16, 16->16
constexpr int LAYER_FLOATS = 2 * MLP48 + MLP16;      // msg, edge, node
constexpr int GNN_CFG_DEFAULT = 6;  // rational-erfc GELU + tensor-core per-edge MLPs

}  // namespace b200spai

struct b200spai_graph_s {
  long long N = 0, E = 0;
  int *perm = nullptr, *src = nullptr, *dst = nullptr, *segptr = nullptr;
  // row-sorted view (GraphSpmv without transpose aggregates at edge_index[0]): position p is original edge
  // rperm[p], its column is rcol[p]; rows own rsegptr[i]..rsegptr[i+1]
  int *rperm = nullptr, *rcol = nullptr, *rsegptr = nullptr;
  float *ef = nullptr, *msg = nullptr, *x0 = nullptr, *x1 = nullptr;  // edge features, messages [E,16], node features
};

struct b200spai_gnn_s {
  int node_in = 0, edge_in = 0, b = 1, bb = 1, L = 4;
  float* blob = nullptr;
  long long blob_len = 0;
  long long launches = 0;
  const float* w_node_enc() const { return blob; }
  const float* w_edge_enc() const { return blob + b200spai::MLP16; }
  const float* w_layer(int l) const { return blob + 2 * b200spai::MLP16 + (long long)l * b200spai::LAYER_FLOATS; }
  const float* w_dec() const { return blob + 2 * b200spai::MLP16 + (long long)L * b200spai::LAYER_FLOATS; }
  static long long expected_len(int L) { return 2 * b200spai::MLP16 + (long long)L * b200spai::LAYER_FLOATS + b200spai::MLP48; }
};
