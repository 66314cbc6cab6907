"""Pins neural_cg/dataset.py (FolderDataset / to_bcoo_components) to the reference.

Run once in the build container (needs /root/reference; torch_geometric etc. come from ref_shim.py):

    python tests/golden/make_golden_dataset.py

Writes three tiny dataset folders with OUR FolderWriter under tests/golden/folders/ (fixed-topology heat-like b=1 with
shared + per-sample features and two right-hand sides per matrix; free-topology Poisson .mtx b=1; free-topology
elasticity .mtx b=3), reads them back with the UNMODIFIED reference FolderDataset (neural_cg/data.py:339-640) and
stores what it returns in tests/golden/dataset_folders.npz.  tests/test_dataset_host.py reads the same folders with our
reader and compares field by field (no /root/reference needed at test time).
"""
from __future__ import annotations

import os
import shutil
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_shim  # noqa: E402

ref_shim.install("/root/reference")
from neural_cg.data import FolderDataset as RefFolderDataset  # noqa: E402  (reference)

from learningsparsepreconditioner4gpu_b200 import datagen as dg  # noqa: E402
from learningsparsepreconditioner4gpu_b200.neural_cg.dataset import FolderWriter  # noqa: E402

FOLDERS = os.path.join(HERE, "folders")
FIELDS = ("x", "edge_index", "edge_attr", "matrix_values", "mask", "diagonal", "inv_diag", "rsqrt_diag")

CASES = {
    "heat_fixed": dict(is_fixed_topology=True, block_size=1, has_shared_features=True, use_node_features=True,
                       use_matrix_as_edge_feature=True, use_mask_as_node_feature=True, use_node_features_as_edge_feature=False,
                       use_edge_features_as_node_feature="disable", use_random_rhs=False, normalize_matrix="mean"),
    "poisson_free": dict(is_fixed_topology=False, block_size=1, has_shared_features=False, use_node_features=True,
                         use_matrix_as_edge_feature=True, use_mask_as_node_feature=True, use_node_features_as_edge_feature=False,
                         use_edge_features_as_node_feature="mean", use_random_rhs=True, normalize_matrix="mean"),
    "elast_free": dict(is_fixed_topology=False, block_size=3, has_shared_features=False, use_node_features=True,
                       use_matrix_as_edge_feature=True, use_mask_as_node_feature=True, use_node_features_as_edge_feature=False,
                       use_edge_features_as_node_feature="disable", use_random_rhs=True, normalize_matrix="mean"),
}


def scipy_of(sys_: dg.GridSystem, b: int):
    from scipy.sparse import bsr_matrix, csr_matrix

    ei = sys_.edge_index.cpu().numpy()
    vals = sys_.values.cpu().numpy().astype(np.float64).reshape(-1, b, b)
    n = sys_.num_nodes
    if b == 1:
        return csr_matrix((vals.reshape(-1), (ei[0], ei[1])), shape=(n, n))
    indptr = np.concatenate([[0], np.cumsum(np.bincount(ei[0], minlength=n))])
    return csr_matrix(bsr_matrix((vals, ei[1], indptr), shape=(n * b, n * b)))


def write_folders():
    shutil.rmtree(FOLDERS, ignore_errors=True)
    # fixed topology: same jittered grid, three coefficient fields
    w = FolderWriter(os.path.join(FOLDERS, "heat_fixed"), block_size=1, is_fixed_topology=True, save_rhs=2)
    base = dg.heat3d(4, 4, 5, seed=1)
    topo = scipy_of(base, 1)
    topo.data[:] = 1.0
    w.write_topology(topo)
    w.write_shared_features(base.xyz.numpy())
    for k in range(3):
        s = dg.heat3d(4, 4, 5, seed=1, var=0.5 + 0.2 * k)
        w.append(scipy_of(s, 1), mask=np.ones((s.num_nodes, 1)), features=s.node_field.numpy(), seed=10 + k)
    # free topology, scalar
    w = FolderWriter(os.path.join(FOLDERS, "poisson_free"), block_size=1, is_fixed_topology=False)
    for k, g in enumerate((4, 5)):
        s = dg.poisson3d(g, seed=20 + k)
        mask = dg.dirichlet_patch_mask(s.num_nodes, s.edge_index, dg.boundary_node_ids(g, g, g), seed=k, rings=1).numpy()
        w.append(scipy_of(s, 1), mask=mask.reshape(-1, 1), features=s.xyz.numpy())
    # free topology, 3x3 blocks
    w = FolderWriter(os.path.join(FOLDERS, "elast_free"), block_size=3, is_fixed_topology=False)
    s = dg.elasticity3d(4, 3, 3, seed=3)
    mask = dg.clamped_ends_mask(4, 3, 3).numpy()
    w.append(scipy_of(s, 3), mask=mask, features=torch.cat([s.node_field, s.xyz], 1).numpy())


def main():
    write_folders()
    out = {}
    for name, kw in CASES.items():
        ds = RefFolderDataset(load_into_memory=False, prefix=os.path.join(FOLDERS, name), **kw)
        out[f"{name}/len"] = np.int64(ds.len())
        out[f"{name}/num_node_features"] = np.int64(ds.num_node_features)
        out[f"{name}/num_edge_features"] = np.int64(ds.num_edge_features)
        for i in range(ds.len()):
            d = ds.get(i)
            for f in FIELDS:
                if hasattr(d, f) and getattr(d, f) is not None:
                    out[f"{name}/{i}/{f}"] = getattr(d, f).numpy()
    np.savez_compressed(os.path.join(HERE, "dataset_folders.npz"), **out)
    print("wrote", len(out), "arrays;", {k: int(out[k]) for k in out if k.endswith("/len")})


if __name__ == "__main__":
    main()
