"""Mint the golden fixtures in tests/golden/*.npz by running the UNMODIFIED reference sources.

Run once in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

The reference has no tests, fixtures or golden vectors of its own, so the fixtures
are outputs of the reference's own code on seeded inputs:

* make_data                         neural_cg/data.py:218-336
* SimpleTrainingWorkspace(...)      neural_cg/workspace.py:27-94   (random init, torch.manual_seed(42))
  .forward -> G [E,b,b];  .inference_step -> scipy CSR
* AATPE                             neural_cg/nn/basic_layers.py:228-261
* to_csr_cpu                        neural_cg/utils/validate.py:22-51   (A and G, masked and unmasked)
* get_pcg_iter_time_scipy / get_cg_iter_time_scipy / get_pcg_scaled_iter_time_scipy
                                    neural_cg/utils/validate.py:163-201,267-345
* generate_spd_sparse_matrix        datagen/synthetic.py:10-27

torch_geometric / lightning / hydra are absent and are replaced by the stand-ins in ref_shim.py (see
its docstring for exactly what is real and what is restated).  The input systems come from this
repo's generators (the reference's own need tetgen / pymathprim / pyssim).
"""
from __future__ import annotations

import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_shim  # noqa: E402

ref_shim.install("/root/reference")

from neural_cg.data import RawData, make_data  # noqa: E402  (reference)
from neural_cg.nn.basic_layers import AATPE  # noqa: E402  (reference)
from neural_cg.utils import validate as rv  # noqa: E402  (reference)
from neural_cg.workspace import SimpleTrainingWorkspace  # noqa: E402  (reference)

sys.path.insert(0, "/root/reference/datagen")
import importlib.util  # noqa: E402

_spec = importlib.util.spec_from_file_location("ref_synthetic", "/root/reference/datagen/synthetic.py")
ref_synth = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(ref_synth)

from learningsparsepreconditioner4gpu_b200 import datagen as dg  # noqa: E402

MLP = {"pre_norm": "none", "hidden_channels": 16, "num_layers": 2}
MLP_LN = {"pre_norm": "layer", "hidden_channels": 16, "num_layers": 2}
GNN_CFG = dict(
    node_encoder=MLP, edge_encoder=MLP, node_decoder=MLP, edge_decoder=MLP, num_mp_layers=4,
    node_residual=True, edge_residual=True, node_features=16, edge_features=16,
    node_mlp=MLP_LN, edge_mlp=MLP_LN, msg_mlp=MLP_LN, msg_norm=True, aggr="add",
)
EPS = 3.0e-3


def build_workspace(f_node, f_edge, block_size, seed=42):
    torch.manual_seed(seed)
    ws = SimpleTrainingWorkspace(
        batch_size=1, batch_less=False, block_size=block_size, test_max_iter=1000,
        optimizer=SimpleNamespace(name="adamw", params={}), scheduler=SimpleNamespace(name="exp", params={}),
        loss=SimpleNamespace(name="RelativeL2Loss_ANorm", params=None), inspect_norms=False,
        check_converge=False, check_methods=[], check_devices=[],
        node_features=f_node, edge_features=f_edge, gnn=GNN_CFG, epsilon=EPS,
    )
    ws.eval()
    return ws


def csr_fields(prefix, m):
    return {prefix + "_indptr": m.indptr.astype(np.int32), prefix + "_indices": m.indices.astype(np.int32),
            prefix + "_data": m.data.astype(np.float64)}


def run_case(name, edge_index, values, mask, node_features, block_size, feat_kwargs, seed=42):
    ei = edge_index.numpy().astype(np.int64)
    vals = values.numpy().astype(np.float64).reshape(-1, block_size, block_size)
    n_nodes = mask.shape[0]
    mask = np.asarray(mask, dtype=np.float64).reshape(n_nodes, block_size)
    # diagonal blocks -> scalar diagonal
    diag = np.zeros((n_nodes, block_size))
    on_diag = ei[0] == ei[1]
    diag[ei[0][on_diag]] = np.diagonal(vals[on_diag], axis1=1, axis2=2)
    raw = RawData(vals, diag, ei, None if node_features is None else np.asarray(node_features, dtype=np.float64),
                  None, None, mask, n_nodes, block_size)
    torch.manual_seed(seed + 1)
    data = make_data(raw, is_inference=False, use_random_rhs=True, normalize_matrix="mean",
                     use_matrix_as_edge_feature=True, use_mask_as_node_feature=True, **feat_kwargs)
    ws = build_workspace(data.x.shape[1], data.edge_attr.shape[1], block_size, seed)
    sd = {k: v.detach().numpy().copy() for k, v in ws.state_dict().items()}
    with torch.no_grad():
        G = ws(data.x, data.edge_index, data.edge_attr)
        G_csr_step, _ = ws.inference_step(data)
        xr = torch.randn(n_nodes, block_size, generator=torch.Generator().manual_seed(7))
        aat = AATPE(EPS)(xr, data.edge_index, G, mask=data.mask)
        aat_diag = AATPE(EPS)(xr, data.edge_index, G, mask=data.mask, diag=data.inv_diag)
    n = n_nodes * block_size
    A_csr = rv.to_csr_cpu(data.edge_index, data.matrix_values, n, data.mask)
    A_csr_nomask = rv.to_csr_cpu(data.edge_index, data.matrix_values, n, None)
    G_csr = rv.to_csr_cpu(data.edge_index, G, n, data.mask)
    assert (G_csr != G_csr_step).nnz == 0
    gt = rv.to_numpy(data.mask).flatten().astype(np.float64)  # rhs="mask" (infer.py:297-299)
    it_pcg = rv.get_pcg_iter_time_scipy(A_csr, gt, G_csr, EPS, rtol=1e-6)
    it_cg = rv.get_cg_iter_time_scipy(A_csr, gt, rtol=1e-6)
    it_scaled = rv.get_pcg_scaled_iter_time_scipy(A_csr, gt, G_csr, EPS, rtol=1e-6)
    it_pcg_1e10 = rv.get_pcg_iter_time_scipy(A_csr, gt, G_csr, EPS, rtol=1e-10)
    # true-residual history of the same scipy solve (same operator the reference builds, validate.py:173-182)
    from oracle.pcg import pcg_scipy

    it2, true_hist, xsol = pcg_scipy(A_csr, gt, G_csr, EPS, rtol=1e-6, history=True)
    assert it2 == it_pcg
    out = dict(
        block_size=np.int64(block_size), epsilon=np.float64(EPS),
        raw_values=vals, raw_edge_index=ei, raw_mask=mask, raw_diag=diag,
        raw_node_features=np.zeros((n_nodes, 0)) if node_features is None else np.asarray(node_features, dtype=np.float64),
        feat_edge_as_node=np.array(feat_kwargs.get("use_edge_features_as_node_feature", "disable")),
        x=data.x.numpy(), edge_attr=data.edge_attr.numpy(), matrix_values=data.matrix_values.numpy(),
        mask=data.mask.numpy(), inv_diag=data.inv_diag.numpy(), rsqrt_diag=data.rsqrt_diag.numpy(),
        diagonal=data.diagonal.numpy(),
        G=G.numpy(), aatpe_in=xr.numpy(), aatpe_out=aat.numpy(), aatpe_diag_out=aat_diag.numpy(),
        iters_pcg=np.int64(it_pcg), iters_cg=np.int64(it_cg), iters_pcg_scaled=np.int64(it_scaled),
        iters_pcg_1e10=np.int64(it_pcg_1e10), true_res_hist=np.asarray(true_hist), x_solution=xsol,
        sd_keys=np.array(list(sd.keys())),
    )
    for k, v in sd.items():
        out["sd/" + k] = v
    out.update(csr_fields("A", A_csr))
    out.update(csr_fields("A_nomask", A_csr_nomask))
    out.update(csr_fields("G", G_csr))
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **out)
    print(f"{name}: N={n_nodes} E={ei.shape[1]} b={block_size} F_in={data.x.shape[1]} F_e={data.edge_attr.shape[1]} "
          f"nnzA={A_csr.nnz} nnzG={G_csr.nnz} iters pcg={it_pcg} cg={it_cg} scaled={it_scaled} "
          f"|G| max={np.abs(G.numpy()).max():.4f} -> {os.path.getsize(path)/1024:.0f} KiB")


def main():
    plain = dict(use_node_features_as_edge_feature=False, use_edge_features_as_node_feature="disable")
    # C2-like: jittered Poisson tet grid, Dirichlet patch
    s = dg.poisson3d(7, 6, 5, jitter=0.2, seed=1)
    m = dg.dirichlet_patch_mask(s.num_nodes, s.edge_index, dg.boundary_node_ids(7, 6, 5), seed=42).numpy()
    run_case("poisson_b1", s.edge_index, s.values, m, s.xyz.numpy(), 1, plain)
    # un-jittered: 8 of 15 stencil entries are exact zeros -> exercises zero pruning in to_csr_cpu
    s = dg.poisson3d(5, jitter=0.0)
    m = dg.dirichlet_patch_mask(s.num_nodes, s.edge_index, dg.boundary_node_ids(5, 5, 5), seed=3).numpy()
    run_case("poisson_zeros_b1", s.edge_index, s.values, m, s.xyz.numpy(), 1, plain)
    # C1-like: heat, coefficient field + xyz + mask, all-ones mask
    s = dg.heat3d(6, jitter=0.2, seed=42)
    feats = torch.cat([s.node_field, s.xyz], dim=1).numpy()
    run_case("heat_b1", s.edge_index, s.values, np.ones(s.num_nodes), feats, 1, plain)
    # C3-like: 3x3-block elasticity, clamped ends, per-DOF mask
    s = dg.elasticity3d(6, 3, 3, seed=5)
    feats = torch.cat([s.node_field, s.xyz], dim=1).numpy()
    run_case("elast_b3", s.edge_index, s.values, dg.clamped_ends_mask(6, 3, 3).numpy(), feats, 3, plain)
    # C4: the reference's own synthetic SPD generator, edge->node mean feature
    A = ref_synth.generate_spd_sparse_matrix(400, sparsity=0.008, condition_amplifier=1e-5,
                                             random_state=np.random.RandomState(42))
    A_mine = dg.generate_spd_sparse_matrix(400, sparsity=0.008, condition_amplifier=1e-5,
                                           random_state=np.random.RandomState(42))
    assert abs(A - A_mine).max() == 0
    ei, v = dg.scipy_to_edges(A)
    run_case("synthetic_b1", ei, v, np.ones(400), None, 1,
             dict(use_node_features_as_edge_feature=False, use_edge_features_as_node_feature="mean"))
    # same generator, mild conditioning, so that an irregular-row system also converges before max_iter
    A = ref_synth.generate_spd_sparse_matrix(300, sparsity=0.01, condition_amplifier=0.3,
                                             random_state=np.random.RandomState(7))
    ei, v = dg.scipy_to_edges(A)
    run_case("synthetic_mild_b1", ei, v, np.ones(300), None, 1,
             dict(use_node_features_as_edge_feature=False, use_edge_features_as_node_feature="mean"))


if __name__ == "__main__":
    main()
