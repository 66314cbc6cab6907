"""CPU check of the GNN backward kernels (csrc/gnn_bwd.cuh) and their driver (csrc/gnn_train.cuh) -- and, since the same
emulation runs them, of the FFMA forward kernels and GraphSpmv (csrc/gnn_kernels.cuh) against the oracle.

The kernels are compiled for the HOST by tests/host_emu (one host thread per CUDA thread, barriers for __syncwarp /
__syncthreads, see cuda_emu.h) and run on the tape a torch forward of the oracle produces; the parameter gradients
are compared with torch autograd through the same oracle (oracle/gnn.py = the reference's NodeEdgeProcessing restated,
neural_cg/nn/gnns.py:77-97, basic_layers.py:103-109,193-225).  This pins the arithmetic, the indexing, the shared-memory
protocol and the buffer wiring of the training path without a GPU; tests/test_training_gpu.py runs the same comparison
through libb200spai.so on the device.

Tolerance: fp32 sums over E edges in a different order than torch's -- 2e-4 of max|grad| per tensor (measured ~1e-6)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from learningsparsepreconditioner4gpu_b200 import systems  # noqa: E402
from learningsparsepreconditioner4gpu_b200.neural_cg.nn import gnns  # noqa: E402
from oracle import gnn as ognn  # noqa: E402

EMU_DIR = os.path.join(ROOT, "tests", "host_emu")
CSRC = os.path.join(ROOT, "learningsparsepreconditioner4gpu_b200", "csrc")
EMU_SO = os.path.join(EMU_DIR, "_gnn_bwd_host.so")


@pytest.fixture(scope="module")
def emu():
    srcs = [os.path.join(EMU_DIR, f) for f in ("gnn_bwd_host.cpp", "cuda_emu.h")] + \
           [os.path.join(CSRC, f) for f in ("gnn_bwd.cuh", "gnn_train.cuh", "gnn_kernels.cuh", "gnn_types.h")]
    if not os.path.exists(EMU_SO) or os.path.getmtime(EMU_SO) < max(os.path.getmtime(s) for s in srcs):
        subprocess.run(["g++", "-O1", "-std=c++20", "-pthread", "-shared", "-fPIC", "-I", CSRC, srcs[0], "-o", EMU_SO], check=True)
    lib = C.CDLL(EMU_SO)
    vp = C.c_void_p
    lib.emu_gnn_backward.argtypes = [C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_longlong] + [vp] * 12 + [C.c_int, vp]
    lib.emu_gnn_backward.restype = C.c_int
    lib.emu_graph_outer.argtypes = [C.c_longlong, C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, vp]
    lib.emu_graph_outer.restype = None
    lib.emu_gnn_forward.argtypes = [C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_longlong] + [vp] * 6 + [C.c_int, C.c_int, vp, vp]
    lib.emu_gnn_forward.restype = C.c_int
    lib.emu_graph_spmv.argtypes = [C.c_longlong, C.c_int, vp, vp, vp, vp, vp, vp, C.c_int, vp]
    lib.emu_graph_spmv.restype = None
    return lib


def graph_arrays(edge_index, N):
    """What b200spai_graph_create builds (csrc/gnn.cu): target-sorted positions and the row-sorted view."""
    row, col = edge_index[0].numpy(), edge_index[1].numpy()
    perm = np.argsort(col, kind="stable").astype(np.int32)
    dst, src = col[perm].astype(np.int32), row[perm].astype(np.int32)
    segptr = np.searchsorted(dst, np.arange(N + 1)).astype(np.int32)
    rperm = np.argsort(row, kind="stable").astype(np.int32)
    rsegptr = np.searchsorted(row[rperm], np.arange(N + 1)).astype(np.int32)
    return perm, src, dst, segptr, rperm, rsegptr


def tile_edges(e_pos):
    """[E,16] by position -> the forward kernels' tiled layout [tile][4 chunks][32 lanes][4 floats]."""
    E = e_pos.shape[0]
    Epad = (E + 31) // 32 * 32
    buf = np.zeros((Epad, 16), dtype=np.float32)
    buf[:E] = e_pos
    return np.ascontiguousarray(buf.reshape(Epad // 32, 32, 4, 4).transpose(0, 2, 1, 3)).reshape(-1)


def oracle_forward_tape(sd, x, edge_index, edge_attr, cfg=None):
    """oracle.gnn.node_edge_processing written out layer by layer so that the per-layer tensors can be kept."""
    cfg = cfg or ognn.DEFAULT_GNN_CFG
    src, dst = edge_index[0], edge_index[1]
    node = ognn.feed_forward(sd, "node_enc.", x, cfg["node_encoder"])
    edge = ognn.feed_forward(sd, "edge_enc.", edge_attr, cfg["edge_encoder"])
    xs, es, aggs = [node], [], []
    for k in range(cfg["num_mp_layers"]):
        p = f"mp_layers.{k}."
        feat = torch.cat([node[dst], node[src], edge], dim=-1)
        agg = ognn._scatter_sum(ognn.feed_forward(sd, p + "msg_mlp.", feat, cfg["msg_mlp"]), dst, node.shape[0])
        node_new = node + ognn.feed_forward(sd, p + "node_mlp.", agg, cfg["node_mlp"])
        edge = edge + ognn.feed_forward(sd, p + "edge_mlp.", feat, cfg["edge_mlp"])
        node = node_new
        xs.append(node)
        es.append(edge)
        aggs.append(agg)
    out = ognn.feed_forward(sd, "edge_dec.", torch.cat([edge, node[src], node[dst]], dim=-1), cfg["edge_decoder"])
    return out, xs, es, aggs


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


CASES = [("poisson5", lambda: (systems.poisson_sample(5, device="cpu"), 1)),
         ("elast_b3", lambda: (systems.elasticity_sample(4, 3, 3, device="cpu"), 3)),
         ("heat", lambda: (systems.heat_sample(4, device="cpu"), 1))]


@pytest.mark.parametrize("name,make", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("sm_count", [1, 3])
def test_backward_kernels_match_autograd(emu, name, make, sm_count):
    s, b = make()
    ws = systems.default_workspace(s, block_size=b)
    torch.manual_seed(7)
    sd = {k: v.detach().clone() for k, v in ws.gnn.state_dict().items()}
    for k, v in sd.items():            # non-trivial LayerNorm affines (random init leaves them at 1 / 0)
        if "pre_norm" in k:
            v.add_(0.3 * torch.randn_like(v))
    params = {k: v.clone().requires_grad_(True) for k, v in sd.items() if "node_msg_norm" not in k}
    N, E, L = s.num_nodes, s.num_edges, 4
    out, xs, es, aggs = oracle_forward_tape(params, s.x, s.edge_index, s.edge_attr)
    with torch.no_grad():              # the tape forward IS the oracle's forward
        _, ref_out = ognn.node_edge_processing({"gnn." + k: v for k, v in sd.items()}, s.x, s.edge_index, s.edge_attr)
    assert torch.equal(out.detach(), ref_out)
    gG = torch.randn(E, b * b)
    out.backward(gG)

    perm, src, dst, segptr, rperm, rsegptr = graph_arrays(s.edge_index, N)
    blob = gnns.pack_state_dict(sd, L)
    xs_np = np.ascontiguousarray(torch.stack([t.detach() for t in xs]).numpy(), dtype=np.float32)
    es_np = np.concatenate([tile_edges(t.detach().numpy()[perm]) for t in es]).astype(np.float32)
    aggs_np = np.ascontiguousarray(torch.stack([t.detach() for t in aggs]).numpy(), dtype=np.float32)
    x_np = np.ascontiguousarray(s.x.numpy(), dtype=np.float32)
    ea_np = np.ascontiguousarray(s.edge_attr.numpy(), dtype=np.float32)
    gG_np = np.ascontiguousarray(gG.numpy(), dtype=np.float32)
    grad = np.full(blob.size, np.nan, dtype=np.float32)
    rc = emu.emu_gnn_backward(N, E, L, s.x.shape[1], s.edge_attr.shape[1], b * b, _ptr(blob), blob.size, _ptr(perm), _ptr(src),
                              _ptr(dst), _ptr(segptr), _ptr(rperm), _ptr(rsegptr), _ptr(xs_np), _ptr(es_np), _ptr(aggs_np),
                              _ptr(x_np), _ptr(ea_np), _ptr(gG_np), sm_count, _ptr(grad))
    assert rc == 0
    assert np.isfinite(grad).all()
    got = gnns.unpack_grad_blob(torch.from_numpy(grad), sd, L)
    assert set(got) == set(params)
    worst = 0.0
    for k, p in params.items():
        scale = max(float(p.grad.abs().max()), 1e-6)
        err = float((got[k] - p.grad).abs().max()) / scale
        worst = max(worst, err)
        assert err < 2e-4, f"{name}: {k} relative error {err:.3e} (max|grad| {scale:.3e})"
    print(f"{name} sm={sm_count}: N={N} E={E} b={b}  worst |dgrad|/max|grad| over {len(params)} tensors = {worst:.2e}")


@pytest.mark.parametrize("b", [1, 3])
@pytest.mark.parametrize("transpose", [0, 1])
def test_graph_outer_matches_autograd(emu, b, transpose):
    """dA of GraphSpmv (basic_layers.py:112-142) against autograd through the oracle's graph_spmv."""
    s, _ = (systems.poisson_sample(4, device="cpu"), 1) if b == 1 else (systems.elasticity_sample(3, 3, 3, device="cpu"), 3)
    N, E = s.num_nodes, s.num_edges
    torch.manual_seed(3)
    A = torch.randn(E, b, b, requires_grad=True)
    X = torch.randn(N, b)
    gy = torch.randn(N, b)
    ognn.graph_spmv(X, s.edge_index, A, None, transpose=bool(transpose)).backward(gy)
    perm, src, dst, *_ = graph_arrays(s.edge_index, N)
    gA = np.zeros((E, b, b), dtype=np.float32)
    emu.emu_graph_outer(E, b, _ptr(perm), _ptr(src), _ptr(dst), _ptr(np.ascontiguousarray(gy.numpy())),
                        _ptr(np.ascontiguousarray(X.numpy())), transpose, 0, _ptr(gA))
    assert np.allclose(gA, A.grad.numpy(), rtol=1e-6, atol=1e-7)
    emu.emu_graph_outer(E, b, _ptr(perm), _ptr(src), _ptr(dst), _ptr(np.ascontiguousarray(gy.numpy())),
                        _ptr(np.ascontiguousarray(X.numpy())), transpose, 1, _ptr(gA))
    assert np.allclose(gA, 2 * A.grad.numpy(), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("name", ["poisson_pair_b1", "elast_b3"])
def test_backward_kernels_reproduce_reference_training_gradients(emu, name):
    """The whole chain against the UNMODIFIED reference: tests/golden/train_<name>.npz holds the loss and every parameter
    gradient of the reference's workspace._shared_step + RelativeL2Loss_ANorm + loss.backward().  Here dL/dG comes from
    autograd through the oracle's AATPE + loss only; the GNN part of the backward is done by the host-compiled CUDA
    kernels and must land on the reference's gradients."""
    from conftest import load_train_case
    from oracle import train as otrain

    batch, sd_full, t = load_train_case(name)
    b, L = int(t["block_size"]), 4
    sd = {k[len("gnn."):]: v for k, v in sd_full.items() if k.startswith("gnn.")}
    N, E = batch["x"].shape[0], batch["edge_index"].shape[1]
    with torch.no_grad():
        out, xs, es, aggs = oracle_forward_tape(sd, batch["x"], batch["edge_index"], batch["edge_attr"])
    G = out.reshape(-1, b, b).clone().requires_grad_(True)
    d = ognn.aatpe(batch["residual"], batch["edge_index"], G, float(t["epsilon"]), batch["mask"])
    loss = otrain.loss_value("RelativeL2Loss_ANorm", batch, d, [int(v) for v in batch["ptr"]])
    loss.backward()
    assert float(loss.detach()) == float(np.float32(t["simple/loss"]))

    perm, src, dst, segptr, rperm, rsegptr = graph_arrays(batch["edge_index"], N)
    blob = gnns.pack_state_dict(sd, L)
    xs_np = np.ascontiguousarray(torch.stack(xs).numpy(), dtype=np.float32)
    es_np = np.concatenate([tile_edges(e.numpy()[perm]) for e in es]).astype(np.float32)
    aggs_np = np.ascontiguousarray(torch.stack(aggs).numpy(), dtype=np.float32)
    x_np = np.ascontiguousarray(batch["x"].numpy(), dtype=np.float32)
    ea_np = np.ascontiguousarray(batch["edge_attr"].numpy(), dtype=np.float32)
    gG_np = np.ascontiguousarray(G.grad.reshape(E, b * b).numpy(), dtype=np.float32)
    grad = np.full(blob.size, np.nan, dtype=np.float32)
    rc = emu.emu_gnn_backward(N, E, L, x_np.shape[1], ea_np.shape[1], b * b, _ptr(blob), blob.size, _ptr(perm), _ptr(src),
                              _ptr(dst), _ptr(segptr), _ptr(rperm), _ptr(rsegptr), _ptr(xs_np), _ptr(es_np), _ptr(aggs_np),
                              _ptr(x_np), _ptr(ea_np), _ptr(gG_np), 2, _ptr(grad))
    assert rc == 0
    got = gnns.unpack_grad_blob(torch.from_numpy(grad), sd, L)
    worst = 0.0
    for k, g in got.items():
        ref = t["simple/grad/gnn." + k]
        err = float(np.abs(g.numpy() - ref).max()) / max(float(np.abs(ref).max()), 1e-12)
        worst = max(worst, err)
        assert err < 2e-4, (k, err)
    print(f"{name}: worst |dgrad|/max|grad| against the reference = {worst:.2e} over {len(got)} tensors")



@pytest.mark.parametrize("name,make", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("gelu_cfg", [0, 1, 2, 3], ids=["ffma_erff", "ffma_rational_erfc", "mma_erff", "mma_rational_erfc_default"])
def test_forward_kernels_match_oracle(emu, name, make, gelu_cfg):
    """The forward kernels compiled for the host against the oracle's NodeEdgeProcessing: G within 1e-5 of max|G| (the bar
    of the GPU tests), final node features likewise.  FFMA variants (B200SPAI_GNN_CFG = 0 / 2): node encoder, edge kernel
    incl. the fused edge encoder, node kernel, decoder.  Tensor-core variants (4 / 6 = the GPU default, csrc/gnn_mma.cuh):
    the same source with `mma.sync.m16n8k8.tf32` done in software on the PTX fragment layout -- the 3xTF32 split, the
    per-lane B-fragment tables, the C-fragment -> A-fragment chaining, the paired first Linear, the LayerNorm quad shuffles
    and the packed GELU all run as written."""
    s, b = make()
    ws = systems.default_workspace(s, block_size=b)
    sd = {k: v.detach().clone() for k, v in ws.gnn.state_dict().items()}
    gen = torch.Generator().manual_seed(5)
    for k, v in sd.items():
        if "pre_norm" in k:
            v.add_(0.3 * torch.randn(v.shape, generator=gen))
    with torch.no_grad():
        node_ref, out_ref = ognn.node_edge_processing({"gnn." + k: v for k, v in sd.items()}, s.x, s.edge_index, s.edge_attr)
    N, E, L = s.num_nodes, s.num_edges, 4
    perm, src, dst, segptr, _, _ = graph_arrays(s.edge_index, N)
    blob = gnns.pack_state_dict(sd, L)
    x_np = np.ascontiguousarray(s.x.numpy(), dtype=np.float32)
    ea_np = np.ascontiguousarray(s.edge_attr.numpy(), dtype=np.float32)
    g_out = np.full((E, b * b), np.nan, dtype=np.float32)
    node_out = np.full((N, 16), np.nan, dtype=np.float32)
    rc = emu.emu_gnn_forward(N, E, L, x_np.shape[1], ea_np.shape[1], b * b, _ptr(blob), blob.size, _ptr(perm), _ptr(src), _ptr(dst),
                             _ptr(segptr), _ptr(x_np), _ptr(ea_np), gelu_cfg, 2, _ptr(g_out), _ptr(node_out))
    assert rc == 0
    err_g = np.abs(g_out - out_ref.numpy()).max() / np.abs(out_ref.numpy()).max()
    err_x = np.abs(node_out - node_ref.numpy()).max() / np.abs(node_ref.numpy()).max()
    print(f"{name} gelu_cfg={gelu_cfg}: |dG|/max|G| = {err_g:.2e}, |dx|/max|x| = {err_x:.2e}")
    assert err_g <= 1e-5 and err_x <= 1e-5


@pytest.mark.parametrize("b", [1, 3])
@pytest.mark.parametrize("transpose", [0, 1])
@pytest.mark.parametrize("masked", [False, True])
def test_graph_spmv_kernel_matches_oracle(emu, b, transpose, masked):
    """graph_spmv_f32_kernel (GraphSpmv, basic_layers.py:112-142) against the oracle's operator: the kernel adds a target's
    edges in ascending edge order like the CPU scatter_add_ of the reference."""
    s, _ = (systems.poisson_sample(5, device="cpu"), 1) if b == 1 else (systems.elasticity_sample(4, 3, 3, device="cpu"), 3)
    N, E = s.num_nodes, s.num_edges
    gen = torch.Generator().manual_seed(8)
    A = torch.randn(E, b, b, generator=gen)
    X = torch.randn(N, b, generator=gen)
    mask = (torch.rand(N, b, generator=gen) > 0.3).float() if masked else None
    ref = ognn.graph_spmv(X, s.edge_index, A, mask, transpose=bool(transpose)).numpy()
    perm, src, dst, segptr, rperm, rsegptr = graph_arrays(s.edge_index, N)
    rcol = s.edge_index[1].numpy()[rperm].astype(np.int32)
    y = np.full((N, b), np.nan, dtype=np.float32)
    seg, pm, other = (segptr, perm, src) if transpose else (rsegptr, rperm, rcol)
    m_np = None if mask is None else np.ascontiguousarray(mask.numpy())
    emu.emu_graph_spmv(N, b, _ptr(seg), _ptr(pm), _ptr(other), _ptr(np.ascontiguousarray(A.numpy())), _ptr(np.ascontiguousarray(X.numpy())),
                       _ptr(m_np) if m_np is not None else None, transpose, _ptr(y))
    assert np.allclose(y, ref, rtol=1e-6, atol=1e-6)
