// TEST INFRASTRUCTURE (not product): the GNN backward kernels and their driver compiled for the host (cuda_emu.h) and
// exported to tests/test_gnn_bwd_host.py, which compares the result with torch autograd through oracle/gnn.py.
//
//   g++ -O1 -std=c++20 -pthread -shared -fPIC -I <csrc> gnn_bwd_host.cpp -o _gnn_bwd_host.so
#include "cuda_emu.h"

#include "gnn_types.h"
typedef b200spai_graph_s* b200spai_graph_t;
typedef b200spai_gnn_s* b200spai_gnn_t;
#include "gnn_kernels.cuh"
#include "gnn_bwd.cuh"
#include "gnn_train.cuh"

extern "C" {

// All arrays are host arrays laid out like the device arrays of the library:
//   graph: target-sorted positions (perm / src / dst / segptr) and the row-sorted view (rperm / rsegptr);
//   tape:  xs [L+1][N][16], es [L][Epad][16] TILED ([tile][4 chunks][32 lanes][4 floats]), aggs [L][N][16].
int emu_gnn_backward(long long N, long long E, int L, int node_in, int edge_in, int bb, const float* blob, long long blob_len,
                     const int* perm, const int* src, const int* dst, const int* segptr, const int* rperm, const int* rsegptr,
                     const float* xs, const float* es, const float* aggs, const float* x, const float* edge_attr,
                     const float* gG, int sm_count, float* grad_out) {
  emu_sm_count = sm_count;
  b200spai_graph_s g;
  g.N = N;
  g.E = E;
  g.perm = const_cast<int*>(perm);
  g.src = const_cast<int*>(src);
  g.dst = const_cast<int*>(dst);
  g.segptr = const_cast<int*>(segptr);
  g.rperm = const_cast<int*>(rperm);
  g.rsegptr = const_cast<int*>(rsegptr);
  b200spai_gnn_s m;
  m.node_in = node_in;
  m.edge_in = edge_in;
  m.bb = bb;
  m.L = L;
  m.blob = const_cast<float*>(blob);
  m.blob_len = blob_len;
  if (blob_len != b200spai_gnn_s::expected_len(L)) return -2;
  b200spai_gnn_tape_s t;
  t.N = N;
  t.E = E;
  t.Epad = ((E + 31) / 32) * 32;
  t.L = L;
  t.max_ctas = sm_count * 4;
  const size_t xsz = (size_t)N * 16, esz = (size_t)t.Epad * 16;
  t.xs = const_cast<float*>(xs);
  t.es = const_cast<float*>(es);
  t.aggs = const_cast<float*>(aggs);
  const float nan = std::nanf("");
  std::vector<float> ge(esz, nan), gxs(esz, nan), gxd(esz, nan), gx(xsz, nan), gagg(xsz, nan),
      partial((size_t)t.max_ctas * BWD_NW_MAX, nan);
  t.ge = ge.data();
  t.gxs = gxs.data();
  t.gxd = gxd.data();
  t.gx = gx.data();
  t.gagg = gagg.data();
  t.partial = partial.data();
  t.filled = true;
  return gnn_backward_impl(&m, &g, &t, x, edge_attr, gG, grad_out, nullptr);
}

// gradient of GraphSpmv w.r.t. the matrix values
void emu_graph_outer(long long E, int b, const int* perm, const int* src, const int* dst, const float* gy, const float* x,
                     int transpose, int accumulate, float* gA) {
  B2_LAUNCH(b200spai::graph_outer_f32_kernel, (E + 255) / 256, 256, 0, nullptr, src, dst, perm, E, b, gy, x, transpose, accumulate, gA);
}

}  // extern "C"
