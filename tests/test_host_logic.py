"""CPU tests of the host side: the neural_cg mirror (parameter containers, weight packing, featurisation,
batching) and the C-ABI library surface.  No compute call is made (there is no GPU here)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, golden_state_dict, load_golden

PKG = "learningsparsepreconditioner4gpu_b200"
MLP = {"pre_norm": "none", "hidden_channels": 16, "num_layers": 2}
MLP_LN = {"pre_norm": "layer", "hidden_channels": 16, "num_layers": 2}
GNN_CFG = dict(node_encoder=MLP, edge_encoder=MLP, node_decoder=MLP, edge_decoder=MLP, num_mp_layers=4,
               node_residual=True, edge_residual=True, node_features=16, edge_features=16,
               node_mlp=MLP_LN, edge_mlp=MLP_LN, msg_mlp=MLP_LN, msg_norm=True, aggr="add")


def make_workspace(f_node, f_edge, b, seed=42, scaled=False, **over):
    from types import SimpleNamespace

    from learningsparsepreconditioner4gpu_b200.neural_cg.scaled_workspace import ScaledTrainingWorkspace
    from learningsparsepreconditioner4gpu_b200.neural_cg.workspace import SimpleTrainingWorkspace

    torch.manual_seed(seed)
    kw = dict(batch_size=1, batch_less=False, block_size=b, test_max_iter=1000,
              optimizer=SimpleNamespace(name="adamw", params={}), scheduler=SimpleNamespace(name="exp", params={}),
              loss=SimpleNamespace(name="RelativeL2Loss_ANorm", params=None), inspect_norms=False,
              check_converge=True, check_methods=["none", "ic", "diagonal"], check_devices=["cuda"],
              node_features=f_node, edge_features=f_edge, gnn=GNN_CFG, epsilon=3.0e-3)
    kw.update(over)
    return (ScaledTrainingWorkspace if scaled else SimpleTrainingWorkspace)(**kw).eval()


def test_library_exports_every_declared_symbol():
    from learningsparsepreconditioner4gpu_b200 import _lib

    header = open(os.path.join(ROOT, "include", "b200spai.h")).read()
    declared = set(re.findall(r"\b(b200spai_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 25
    lib = ctypes.CDLL(_lib.LIB_PATH) if os.path.exists(_lib.LIB_PATH) else _lib.load()
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert declared == set(_lib.SIGNATURES) | {"b200spai_last_error"}
    _lib.load()
    assert lib.b200spai_version() >= 100


def test_random_init_reproduces_reference_state_dict(golden):
    """Same module registration order + same torch seed => the weights the reference workspace draws."""
    name, g = golden
    ws = make_workspace(g["x"].shape[1], g["edge_attr"].shape[1], int(g["block_size"]))
    ref = golden_state_dict(g)
    mine = ws.state_dict()
    assert list(mine.keys()) == [str(k) for k in g["sd_keys"]], name
    for k, v in ref.items():
        assert torch.equal(mine[k], v), (name, k)


def test_state_dict_loads_and_packs(golden):
    from learningsparsepreconditioner4gpu_b200.neural_cg.nn import gnns

    name, g = golden
    ws = make_workspace(g["x"].shape[1], g["edge_attr"].shape[1], int(g["block_size"]), seed=1)
    ws.load_state_dict(golden_state_dict(g))
    blob = gnns.pack_state_dict(ws.gnn.state_dict(), 4)
    assert blob.dtype == np.float32 and blob.size == gnns.blob_length(4)
    sd = ws.gnn.state_dict()
    # node encoder, first Linear: blob holds W^T zero-padded to 16 input rows
    W = sd["node_enc.lift.0.weight"].numpy()
    assert np.array_equal(blob[: 16 * 16].reshape(16, 16)[: W.shape[1]], W.T)
    assert not blob[: 16 * 16].reshape(16, 16)[W.shape[1]:].any()
    # LayerNorm folding of layer 0 msg_mlp: W1' = g * W1^T, b1' = b1 + W1 c
    off = 2 * gnns.MLP16
    W1 = sd["mp_layers.0.msg_mlp.lift.0.weight"].double().numpy()
    gam = sd["mp_layers.0.msg_mlp.pre_norm.weight"].double().numpy()
    c = sd["mp_layers.0.msg_mlp.pre_norm.bias"].double().numpy()
    b1 = sd["mp_layers.0.msg_mlp.lift.0.bias"].double().numpy()
    assert np.allclose(blob[off: off + 48 * 16].reshape(48, 16), (W1 * gam).T.astype(np.float32))
    assert np.allclose(blob[off + 48 * 16: off + 48 * 16 + 16], (b1 + W1 @ c).astype(np.float32))


def test_unsupported_architecture_raises():
    from learningsparsepreconditioner4gpu_b200.neural_cg.nn.gnns import NodeEdgeProcessing

    bad = dict(GNN_CFG, msg_mlp={"pre_norm": "layer", "hidden_channels": 32, "num_layers": 2})
    with pytest.raises(NotImplementedError):
        NodeEdgeProcessing(node_in_features=4, node_out_features=None, edge_in_features=1, edge_out_features=1, **bad)
    with pytest.raises(NotImplementedError):
        NodeEdgeProcessing(node_in_features=4, node_out_features=3, edge_in_features=1, edge_out_features=1, **GNN_CFG)


def test_no_cpu_path():
    """The product fails loudly without a CUDA device instead of falling back."""
    from learningsparsepreconditioner4gpu_b200 import _lib
    from learningsparsepreconditioner4gpu_b200.neural_cg.nn.basic_layers import FeedForward
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils import validate

    if torch.cuda.is_available():
        pytest.skip("needs a CPU-only box")
    g = load_golden("poisson_b1")
    ws = make_workspace(g["x"].shape[1], 1, 1)
    with pytest.raises(_lib.B200SpaiError):
        ws(torch.from_numpy(g["x"]), torch.from_numpy(g["raw_edge_index"]), torch.from_numpy(g["edge_attr"]))
    with pytest.raises(_lib.B200SpaiError):
        validate.to_csr_cpu(torch.from_numpy(g["raw_edge_index"]), torch.from_numpy(g["matrix_values"]), 210, None)
    with pytest.raises(_lib.B200SpaiError):
        FeedForward(4, 16, 16, 2)(torch.zeros(3, 4))
    with pytest.raises(ValueError):
        validate.get_pcg_iter_time(None, None, None, 1e-3, device="cpu")


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, PKG)):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), os.path.join(dirpath, f)


def test_make_data_matches_reference(golden):
    from learningsparsepreconditioner4gpu_b200.neural_cg.data import RawData, make_data

    name, g = golden
    nf = g["raw_node_features"]
    b = int(g["block_size"])
    raw = RawData(g["raw_values"], g["raw_diag"], g["raw_edge_index"], None if nf.shape[1] == 0 else nf, None, None,
                  g["raw_mask"], g["raw_mask"].shape[0], b)
    d = make_data(raw, use_edge_features_as_node_feature=str(g["feat_edge_as_node"]))
    for k in ("x", "edge_attr", "matrix_values", "mask", "inv_diag", "rsqrt_diag", "diagonal"):
        assert np.array_equal(d[k].numpy(), g[k]), (name, k)
    assert d.residual.shape == (raw.num_nodes, b)
    # torch-tensor inputs (the on-device featurisation used for the 10 M-DOF systems) agree to fp32 rounding
    raw_t = RawData(torch.from_numpy(g["raw_values"]), torch.from_numpy(g["raw_diag"]), torch.from_numpy(g["raw_edge_index"]),
                    None if nf.shape[1] == 0 else torch.from_numpy(nf), None, None, torch.from_numpy(g["raw_mask"]),
                    raw.num_nodes, b)
    dt = make_data(raw_t, use_edge_features_as_node_feature=str(g["feat_edge_as_node"]), is_inference=True)
    for k in ("x", "edge_attr", "matrix_values"):
        assert np.allclose(dt[k].numpy(), g[k], rtol=2e-7, atol=0), (name, k)


def test_collate_block_diagonal():
    from learningsparsepreconditioner4gpu_b200.neural_cg.data import Data, collate

    a = Data(x=torch.ones(3, 2), mask=torch.ones(3, 1), edge_index=torch.tensor([[0, 1, 2], [0, 1, 2]]), edge_attr=torch.ones(3, 1))
    b = Data(x=torch.zeros(2, 2), mask=torch.ones(2, 1), edge_index=torch.tensor([[0, 1], [1, 0]]), edge_attr=torch.zeros(2, 1))
    c = collate([a, b])
    assert c.ptr.tolist() == [0, 3, 5] and c.batch.tolist() == [0, 0, 0, 1, 1]
    assert c.edge_index.tolist() == [[0, 1, 2, 3, 4], [0, 1, 2, 4, 3]]
    assert c.x.shape == (5, 2) and c.edge_attr.shape == (5, 1)
    assert c.to("cpu").num_nodes == 5


def test_diagonal_of():
    from learningsparsepreconditioner4gpu_b200.neural_cg.data import diagonal_of

    g = load_golden("elast_b3")
    d = diagonal_of(g["raw_edge_index"], g["raw_values"], g["raw_mask"].shape[0])
    assert np.array_equal(d, g["raw_diag"])
    dt = diagonal_of(torch.from_numpy(g["raw_edge_index"]), torch.from_numpy(g["raw_values"]), g["raw_mask"].shape[0])
    assert np.array_equal(dt.numpy(), g["raw_diag"])


def test_batch_solver_needs_cuda():
    from learningsparsepreconditioner4gpu_b200 import _lib
    from learningsparsepreconditioner4gpu_b200.batch import BatchSolver

    if torch.cuda.is_available():
        pytest.skip("needs a CPU-only box")
    with pytest.raises(_lib.B200SpaiError):
        BatchSolver(lambda: None, streams=2)
    from learningsparsepreconditioner4gpu_b200.batch import CollatedBatchSolver

    with pytest.raises(_lib.B200SpaiError):
        CollatedBatchSolver(None)               # the collated (one launch per wave) path has no CPU fallback either


def test_weight_blob_pack_and_gradient_unpack_are_consistent():
    """Training plumbing on the host: pack_state_dict_torch (device-side repack after an optimiser step) gives the bits of
    pack_state_dict, and unpack_grad_blob is the exact adjoint of the packing (LayerNorm affine un-folded, transposes and
    padding undone): for L = <blob(params), c> autograd's dL/dparams must equal unpack_grad_blob(c)."""
    from learningsparsepreconditioner4gpu_b200.neural_cg.nn import gnns

    ws = make_workspace(5, 9, 3)
    sd = {k: v.detach().clone() for k, v in ws.gnn.state_dict().items()}
    gen = torch.Generator().manual_seed(0)
    for k, v in sd.items():
        if "pre_norm" in k:                                   # random init leaves the affines at 1 / 0
            v.copy_(torch.randn(v.shape, generator=gen))
    L = 4
    a = gnns.pack_state_dict(sd, L)
    assert a.size == gnns.blob_length(L) and np.array_equal(a, gnns.pack_state_dict_torch(sd, L).numpy())

    params = {k: v.clone().double().requires_grad_(True) for k, v in sd.items() if "node_msg_norm" not in k}

    def pack_autograd(p):
        parts = []
        for pre, _, k_pad in gnns.blob_layout(L):
            W1, b1 = p[pre + "lift.0.weight"], p[pre + "lift.0.bias"]
            if pre + "pre_norm.weight" in p:
                b1 = b1 + W1 @ p[pre + "pre_norm.bias"]
                W1 = W1 * p[pre + "pre_norm.weight"][None, :]
            W3 = p[pre + "proj.0.weight"]
            W1t = torch.cat([W1.T, W1.new_zeros(k_pad - W1.shape[1], 16)])
            W3t = torch.cat([W3.T, W3.new_zeros(16, 16 - W3.shape[0])], dim=1)
            b3 = torch.cat([p[pre + "proj.0.bias"], W3.new_zeros(16 - W3.shape[0])])
            parts += [W1t.reshape(-1), b1, p[pre + "body.0.0.weight"].T.reshape(-1), p[pre + "body.0.0.bias"], W3t.reshape(-1), b3]
        return torch.cat(parts)

    blob = pack_autograd(params)
    assert np.allclose(blob.detach().numpy(), a, rtol=0, atol=1e-6)
    coef = torch.randn(a.size, dtype=torch.float64, generator=gen)
    (blob * coef).sum().backward()
    got = gnns.unpack_grad_blob(coef, {k: v.detach() for k, v in params.items()}, L)
    assert set(got) == set(params) and len(got) == 114
    for k, v in params.items():
        assert got[k].shape == v.shape and torch.allclose(got[k], v.grad, rtol=0, atol=1e-12), k
