"""The five BASELINE.json configurations as ready-to-run samples (``neural_cg.data.Data``).

Each builder pairs a generator from ``datagen`` with the featurisation the reference's training script
selects for that experiment (SURVEY.md Appendix A "per-experiment overrides"), so that tests, ``smoke()``
and ``bench.py`` all feed the hot path the same inputs:

    C1 heat         heat3d            x = [field, xyz, mask]            (training/heat.sh)
    C2 poisson      poisson3d         x = [xyz, mask], Dirichlet patch  (training/poisson_tetmesh.sh)
    C3 elasticity   elasticity3d      b = 3, x = [deform, xyz, mask3]   (training/elast_twist.sh)
    C4 synthetic    generate_spd...   x = [mask, mean incoming edge]    (training/synthetic.sh)
    C5 poisson10m   poisson3d(216)    as C2, generated and featurised on the GPU

Everything here is input construction (outside every timed region); nothing is computed by the library.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch

from . import datagen as dg
from .neural_cg.data import Data, RawData, diagonal_of, make_data


def _finish(sys_: dg.GridSystem, feats, mask, b: int, device, **feat_kw) -> Data:
    on_dev = sys_.values.is_cuda
    if on_dev:  # large systems: featurise where they were generated
        vals = sys_.values.reshape(-1, b, b)
        ei = sys_.edge_index
        raw = RawData(vals, diagonal_of(ei, vals, sys_.num_nodes), ei, feats, None, None,
                      mask.to(vals.device), sys_.num_nodes, b)
    else:
        vals = sys_.values.numpy().reshape(-1, b, b)
        ei = sys_.edge_index.numpy()
        raw = RawData(vals, diagonal_of(ei, vals, sys_.num_nodes), ei, None if feats is None else feats.numpy(),
                      None, None, mask.numpy().reshape(sys_.num_nodes, -1), sys_.num_nodes, b)
    d = make_data(raw, is_inference=True, **feat_kw)
    d.ptr = torch.tensor([0, sys_.num_nodes], dtype=torch.long, device=d.x.device)
    return d.to(device) if device is not None else d


def poisson_sample(n: int, ny: Optional[int] = None, nz: Optional[int] = None, seed: int = 0, jitter: float = 0.2,
                   gen_device="cpu", device=None, mask_seed: int = 42) -> Data:
    """C2 / C5: jittered Kuhn-tet Poisson, Dirichlet patch grown 2 rings around one boundary node."""
    ny, nz = ny or n, nz or n
    s = dg.poisson3d(n, ny, nz, jitter=jitter, seed=seed, device=gen_device)
    mask = dg.dirichlet_patch_mask(s.num_nodes, s.edge_index, dg.boundary_node_ids(n, ny, nz), seed=mask_seed)
    return _finish(s, s.xyz, mask.reshape(-1, 1), 1, device)


def heat_sample(n: int = 18, ny: Optional[int] = None, nz: Optional[int] = None, seed: int = 42, gen_device="cpu",
                device=None) -> Data:
    """C1: L(kappa) + 1e-4 M, features [field, xyz, mask], mask all ones (18x18x19 = 6156 ~ the 6276-row bunny)."""
    ny, nz = ny or n, nz or (n + 1)
    s = dg.heat3d(n, ny, nz, jitter=0.2, seed=seed, device=gen_device)
    feats = torch.cat([s.node_field, s.xyz], dim=1)
    mask = torch.ones(s.num_nodes, 1, dtype=torch.float64, device=s.values.device)
    return _finish(s, feats, mask, 1, device)


def elasticity_sample(nx: int = 64, ny: int = 16, nz: int = 16, seed: int = 0, gen_device="cpu", device=None) -> Data:
    """C3: 3x3-block linear elasticity + mass on the twist beam box, both end faces clamped."""
    s = dg.elasticity3d(nx, ny, nz, seed=seed, device=gen_device)
    feats = torch.cat([s.node_field, s.xyz], dim=1)
    mask = dg.clamped_ends_mask(nx, ny, nz).to(s.values.device)
    return _finish(s, feats, mask, 3, device)


def synthetic_sample(n: int = 10240, sparsity: float = 3e-4, amplifier: float = 1e-5, seed: int = 42, device=None) -> Data:
    """C4: the reference's synthetic SPD generator (datagen/synthetic.py:10-27), edge->node mean feature."""
    A = dg.generate_spd_sparse_matrix(n, sparsity, amplifier, random_state=np.random.RandomState(seed))
    ei, v = dg.scipy_to_edges(A)
    raw = RawData(v.numpy().reshape(-1, 1, 1), diagonal_of(ei.numpy(), v.numpy().reshape(-1, 1, 1), n), ei.numpy(), None,
                  None, None, np.ones((n, 1)), n, 1)
    d = make_data(raw, is_inference=True, use_edge_features_as_node_feature="mean")
    d.ptr = torch.tensor([0, n], dtype=torch.long)
    return d.to(device) if device is not None else d


SHAPES = ("heat", "elast", "synth10240", "synth15360", "synth20480")


def shape_sample(name: str, device=None):
    """(sample, block_size) of a named BASELINE shape at its stated size: ``heat`` (C1, 18x18x19 = 6156 rows, the size
    class of the 6276-row bunny mesh of datagen/config/heat.yaml:1), ``elast`` (C3, unit_box(64,16,16), b = 3, 49152 DOF,
    datagen/config/elast_twist.yaml:12-13), ``synth<n>`` (C4, datagen/synthetic.py:10-27 verbatim with n rows;
    datagen/config/synthetic.yaml:15-20 draws n in [10240, 20480])."""
    if name == "heat":
        return heat_sample(18, device=device), 1
    if name == "elast":
        return elasticity_sample(64, 16, 16, device=device), 3
    if name.startswith("synth"):
        return synthetic_sample(int(name[5:]), device=device), 1
    raise ValueError(f"unknown shape {name!r}; one of {SHAPES}")


def default_workspace(sample: Data, block_size: int = 1, seed: int = 42, epsilon: float = 3.0e-3, scaled: bool = False):
    """Workspace with the reference's resolved default config and random-init weights."""
    from types import SimpleNamespace

    from .neural_cg.scaled_workspace import ScaledTrainingWorkspace
    from .neural_cg.workspace import SimpleTrainingWorkspace

    mlp = {"pre_norm": "none", "hidden_channels": 16, "num_layers": 2}
    mlp_ln = {"pre_norm": "layer", "hidden_channels": 16, "num_layers": 2}
    gnn = dict(node_encoder=mlp, edge_encoder=mlp, node_decoder=mlp, edge_decoder=mlp, num_mp_layers=4,
               node_residual=True, edge_residual=True, node_features=16, edge_features=16,
               node_mlp=mlp_ln, edge_mlp=mlp_ln, msg_mlp=mlp_ln, msg_norm=True, aggr="add")
    torch.manual_seed(seed)
    cls = ScaledTrainingWorkspace if scaled else SimpleTrainingWorkspace
    return cls(batch_size=1, batch_less=False, block_size=block_size, test_max_iter=1000,
               optimizer=SimpleNamespace(name="adamw", params={}), scheduler=SimpleNamespace(name="exp", params={}),
               loss=SimpleNamespace(name="RelativeL2Loss_ANorm", params=None), inspect_norms=False,
               check_converge=True, check_methods=["none", "diagonal"], check_devices=["cuda"],
               node_features=sample.x.shape[1], edge_features=sample.edge_attr.shape[1], gnn=gnn, epsilon=epsilon).eval()
