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
  // every buffer the kernels write sits between two guard bands holding a canary: an out-of-bounds store of up to GUARD
  // floats on either side is detected after the run (compute-sanitizer is not available on the GPU pool; this and the
  // comparison with the CPU reference are the memory check of these kernels).  Contents start as NaN: a read of a value no
  // kernel produced poisons the gradient.
  constexpr size_t GUARD = 1024;
  const float nan = std::nanf(""), canary = -7.25e30f;
  struct Guarded {
    std::vector<float> v;
    size_t n;
    Guarded(size_t n_, float fill, float can) : v(n_ + 2 * GUARD, fill), n(n_) {
      for (size_t i = 0; i < GUARD; ++i) v[i] = v[GUARD + n + i] = can;
    }
    float* data() { return v.data() + GUARD; }
    bool intact(float can) const {
      for (size_t i = 0; i < GUARD; ++i)
        if (v[i] != can || v[GUARD + n + i] != can) return false;
      return true;
    }
  };
  Guarded ge(esz, nan, canary), gxs(esz, nan, canary), gxd(esz, nan, canary), gx(xsz, nan, canary), gagg(xsz, nan, canary),
      partial((size_t)t.max_ctas * BWD_NW_MAX, nan, canary), grad((size_t)blob_len, nan, canary);
  t.ge = ge.data();
  t.gxs = gxs.data();
  t.gxd = gxd.data();
  t.gx = gx.data();
  t.gagg = gagg.data();
  t.partial = partial.data();
  t.filled = true;
  const int rc = gnn_backward_impl(&m, &g, &t, x, edge_attr, gG, grad.data(), nullptr);
  if (rc) return rc;
  for (const Guarded* b : {&ge, &gxs, &gxd, &gx, &gagg, &partial, &grad})
    if (!b->intact(canary)) return -77;  // a kernel stored outside its buffer
  std::memcpy(grad_out, grad.data(), (size_t)blob_len * sizeof(float));
  return 0;
}

// gradient of GraphSpmv w.r.t. the matrix values
void emu_graph_outer(long long E, int b, const int* perm, const int* src, const int* dst, const float* gy, const float* x,
                     int transpose, int accumulate, float* gA) {
  B2_LAUNCH(b200spai::graph_outer_f32_kernel, (E + 255) / 256, 256, 0, nullptr, src, dst, perm, E, b, gy, x, transpose, accumulate, gA);
}

}  // extern "C"
