"""ORACLE (test infrastructure, not product): restatement of ``make_data`` (neural_cg/data.py:218-336),
the featurisation that defines the hot path's inputs, for the inference-relevant fields.

    matrix_scale = 1 / mean(|block_values|)          (normalize_matrix "mean"/True, :248-250)
                 = 1 / ||block_values||_F            ("frob", :251-253);  1 for "none"/False
    x            = cat[node_features?, mask?, scatter_<reduce>(edge_attr -> edge_index[1])?]  (:270-303)
    edge_attr    = float32(scale * blocks).flatten(1)  (+ optional [x[row], x[col]])          (:282-292)
    matrix_values= float32(blocks * scale)                                                  (:312)
    diagonal, inv_diag = 1/(d+1e-7), rsqrt_diag = 1/sqrt(d+1e-7)  with d = scale*diag        (:316-323)

Pinned against the reference's own make_data in tests/test_oracle_golden.py.
"""
from __future__ import annotations

import numpy as np
import torch


def make_data(
    block_values, edge_index, mask, node_features=None, diagonals=None, normalize_matrix="mean",
    use_matrix_as_edge_feature=True, use_mask_as_node_feature=True,
    use_node_features_as_edge_feature=False, use_edge_features_as_node_feature="disable",
):
    block_values = np.asarray(block_values, dtype=np.float64)
    ei = torch.as_tensor(np.asarray(edge_index), dtype=torch.long)
    if normalize_matrix is True or normalize_matrix == "mean":
        scale = 1.0 / np.mean(np.abs(block_values))
    elif normalize_matrix == "frob":
        scale = 1.0 / np.linalg.norm(block_values)
    elif normalize_matrix in ("none", False):
        scale = 1.0
    else:
        raise ValueError(normalize_matrix)
    mask_t = torch.tensor(np.asarray(mask), dtype=torch.float32)
    n_nodes = mask_t.shape[0]
    nf = []
    if node_features is not None:
        nf.append(torch.tensor(np.asarray(node_features), dtype=torch.float32))
    if use_mask_as_node_feature:
        nf.append(mask_t)
    ef = []
    if use_matrix_as_edge_feature:
        ef.append(torch.tensor(scale * block_values, dtype=torch.float32).flatten(1))
    if use_node_features_as_edge_feature:
        cat = torch.cat(nf, dim=-1)
        ef += [cat[ei[0]], cat[ei[1]]]
    edge_attr = torch.cat(ef, dim=-1)
    if use_edge_features_as_node_feature != "disable":
        red = use_edge_features_as_node_feature
        acc = torch.zeros(n_nodes, edge_attr.shape[1]).index_add_(0, ei[1], edge_attr)
        if red == "mean":
            cnt = torch.zeros(n_nodes).index_add_(0, ei[1], torch.ones(ei.shape[1])).clamp(min=1)
            acc = acc / cnt[:, None]
        elif red not in ("sum", "add"):
            raise NotImplementedError(red)
        nf.append(acc)
    out = {
        "x": torch.cat(nf, dim=-1),
        "mask": mask_t,
        "edge_index": ei,
        "edge_attr": edge_attr,
        "matrix_values": torch.tensor(block_values * scale, dtype=torch.float32),
        "matrix_scale": scale,
    }
    if diagonals is not None:
        d = np.asarray(diagonals) * scale
        out["diagonal"] = torch.tensor(d, dtype=torch.float32)
        out["inv_diag"] = torch.tensor(1.0 / (d + 1e-7), dtype=torch.float32)
        out["rsqrt_diag"] = torch.tensor(1.0 / np.sqrt(d + 1e-7), dtype=torch.float32)
    return out
