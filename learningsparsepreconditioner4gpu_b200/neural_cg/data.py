"""Host-side mirror of the input contract of ``neural_cg/data.py``: ``RawData`` (:173-183), ``make_data``
(:218-336) and a minimal stand-in for the torch_geometric ``Data`` / batch objects the workspaces consume
(SURVEY.md Appendix C).  Featurisation is one-off per sample and stays in torch (any device); no kernel.

``make_data`` accepts numpy arrays (the reference's case; the matrix scale is then computed with numpy in
float64 exactly as the reference does) or torch tensors on any device (the >= 10 M-DOF systems are
generated and featurised on the GPU).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, List, Literal, Optional, Sequence, Union

import numpy as np
import torch

Array = Union[np.ndarray, torch.Tensor]


@dataclass
class RawData:
    block_values: Optional[Array]   # [E, b, b]
    diagonals: Optional[Array]      # [N, b]
    edge_index: Array               # [2, E]
    node_features: Optional[Array]  # [N, k]
    lhs: Optional[Array]
    rhs: Optional[Array]
    mask: Array                     # [N, b]
    num_nodes: int
    block_size: int


class Data:
    """Attribute bag of tensors with ``.to(device)`` -- the subset of torch_geometric.data.Data the hot
    path touches (``x edge_index edge_attr mask matrix_values diagonal inv_diag rsqrt_diag residual ptr batch``)."""

    def __init__(self, **fields: Any):
        self.__dict__.update(fields)

    def keys(self) -> List[str]:
        return [k for k in self.__dict__ if not k.startswith("_")]

    def __contains__(self, key: str) -> bool:
        return key in self.__dict__

    def __getitem__(self, key: str):
        return self.__dict__[key]

    def to(self, device, non_blocking: bool = False) -> "Data":
        return Data(**{k: (v.to(device, non_blocking=non_blocking) if isinstance(v, torch.Tensor) else v)
                       for k, v in self.__dict__.items()})

    def pin_memory(self) -> "Data":
        return Data(**{k: (v.pin_memory() if isinstance(v, torch.Tensor) and not v.is_cuda else v)
                       for k, v in self.__dict__.items()})

    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])

    @property
    def num_edges(self) -> int:
        return int(self.edge_index.shape[1])

    def __repr__(self) -> str:
        parts = [f"{k}={list(v.shape)}" if isinstance(v, torch.Tensor) else f"{k}={v!r}" for k, v in self.__dict__.items()]
        return "Data(" + ", ".join(parts) + ")"


def collate(samples: Sequence[Data]) -> Data:
    """Block-diagonal batching as the PyG ``DataLoader`` does it (utils/datamodule.py:47-60): node/edge
    tensors concatenated on dim 0, ``edge_index`` on dim 1 with cumulative node offsets, plus ``ptr`` / ``batch``."""
    assert len(samples) > 0
    out: Dict[str, Any] = {}
    offsets = np.cumsum([0] + [s.num_nodes for s in samples])
    for k in samples[0].keys():
        v0 = samples[0][k]
        if not isinstance(v0, torch.Tensor):
            out[k] = [s[k] for s in samples]
        elif k == "edge_index":
            out[k] = torch.cat([s[k] + int(o) for s, o in zip(samples, offsets[:-1])], dim=1)
        elif k in ("ptr", "batch"):
            continue
        else:
            out[k] = torch.cat([s[k] for s in samples], dim=0)
    dev = samples[0].x.device
    out["ptr"] = torch.as_tensor(offsets, dtype=torch.long, device=dev)
    out["batch"] = torch.repeat_interleave(torch.arange(len(samples), device=dev),
                                           torch.as_tensor(np.diff(offsets), device=dev))
    return Data(**out)


def _t(a: Array, dtype, device=None) -> torch.Tensor:
    if isinstance(a, torch.Tensor):
        return a.to(dtype=dtype, device=device if device is not None else a.device)
    return torch.tensor(np.asarray(a), dtype=dtype, device=device)


def matrix_scale_of(block_values: Array, normalize_matrix: Union[bool, str], raw: Optional[RawData] = None) -> float:
    """data.py:244-268."""
    if normalize_matrix is True or normalize_matrix == "mean":
        if isinstance(block_values, torch.Tensor):
            return 1.0 / float(block_values.double().abs().mean())
        return float(1.0 / np.mean(np.abs(block_values)))
    if normalize_matrix == "frob":
        if isinstance(block_values, torch.Tensor):
            return 1.0 / float(torch.linalg.vector_norm(block_values.double()))
        return float(1.0 / np.linalg.norm(block_values))
    if normalize_matrix == "l1":
        bv = _t(block_values, torch.float64).abs()
        ei = _t(raw.edge_index, torch.long, bv.device)
        row_sum = torch.zeros(raw.num_nodes, bv.shape[1], dtype=torch.float64, device=bv.device)
        row_sum.index_add_(0, ei[0], bv.sum(-1))
        return 1.0 / (float(row_sum.max()) + 1e-7)
    if normalize_matrix == "none" or normalize_matrix is False:
        return 1.0
    raise ValueError(f"normalize_matrix={normalize_matrix!r} not supported")


def make_data(
    raw_data: RawData,
    use_matrix_as_edge_feature: bool = True,
    use_mask_as_node_feature: bool = True,
    use_node_features_as_edge_feature: bool = False,
    use_edge_features_as_node_feature: Literal["disable", "sum", "mean"] = "disable",
    use_random_rhs: bool = True,
    normalize_matrix: Union[bool, str] = "mean",
    is_inference: bool = False,
) -> Data:
    assert not (use_node_features_as_edge_feature and use_edge_features_as_node_feature != "disable")
    bv = raw_data.block_values
    dev = bv.device if isinstance(bv, torch.Tensor) else None
    edge_index = _t(raw_data.edge_index, torch.long, dev)
    scale = 1.0
    if bv is not None:
        scale = matrix_scale_of(bv, normalize_matrix, raw_data)
    n_nodes, bsize = raw_data.num_nodes, raw_data.block_size
    mask = _t(raw_data.mask, torch.float32, dev).reshape(n_nodes, -1)

    node_features: List[torch.Tensor] = []
    if raw_data.node_features is not None:
        node_features.append(_t(raw_data.node_features, torch.float32, dev))
    if use_mask_as_node_feature:
        node_features.append(mask)

    edge_features: List[torch.Tensor] = []
    scaled64 = None
    if bv is not None:
        scaled64 = (_t(bv, torch.float64, dev) * scale) if isinstance(bv, torch.Tensor) else None
    if use_matrix_as_edge_feature:
        assert bv is not None
        vals32 = scaled64.to(torch.float32) if scaled64 is not None else torch.tensor(scale * np.asarray(bv), dtype=torch.float32)
        edge_features.append(vals32.flatten(1))
    if use_node_features_as_edge_feature:
        nf = torch.cat(node_features, dim=-1)
        edge_features += [nf[edge_index[0]], nf[edge_index[1]]]
    assert edge_features, "No edge feature found."
    edge_input = torch.cat(edge_features, dim=-1)

    if use_edge_features_as_node_feature != "disable":
        red = use_edge_features_as_node_feature
        acc = torch.zeros(n_nodes, edge_input.shape[1], dtype=torch.float32, device=edge_input.device)
        acc.index_add_(0, edge_index[1], edge_input)
        if red == "mean":
            cnt = torch.zeros(n_nodes, dtype=torch.float32, device=edge_input.device)
            cnt.index_add_(0, edge_index[1], torch.ones_like(edge_index[1], dtype=torch.float32))
            acc = acc / cnt.clamp(min=1)[:, None]
        elif red not in ("sum", "add"):
            raise NotImplementedError(f"use_edge_features_as_node_feature={red!r}")
        node_features.append(acc)
    assert node_features, "No node feature found."

    fields: Dict[str, Any] = {"x": torch.cat(node_features, dim=-1), "mask": mask, "edge_index": edge_index,
                              "edge_attr": edge_input}
    if bv is not None:
        mv = scaled64.to(torch.float32) if scaled64 is not None else torch.tensor(np.asarray(bv) * scale, dtype=torch.float32)
        fields["matrix_values"] = mv.reshape(-1, bsize, bsize)
    else:
        assert not is_inference, "Training depends on matrix values."
    if raw_data.diagonals is not None:
        if isinstance(raw_data.diagonals, torch.Tensor):
            d = raw_data.diagonals.double() * scale
            fields["diagonal"] = d.float()
            fields["inv_diag"] = (1.0 / (d + 1e-7)).float()
            fields["rsqrt_diag"] = (1.0 / torch.sqrt(d + 1e-7)).float()
        else:
            d = np.asarray(raw_data.diagonals) * scale
            fields["diagonal"] = torch.tensor(d, dtype=torch.float32)
            fields["inv_diag"] = torch.tensor(1.0 / (d + 1e-7), dtype=torch.float32)
            fields["rsqrt_diag"] = torch.tensor(1.0 / np.sqrt(d + 1e-7), dtype=torch.float32)
    if not is_inference:
        rhs = torch.randn(n_nodes, bsize, dtype=torch.float32).to(mask.device)
        if not use_random_rhs:
            assert raw_data.rhs is not None
            rhs = _t(raw_data.rhs, torch.float32, dev)
            if raw_data.lhs is not None:
                fields["gt"] = _t(raw_data.lhs, torch.float32, dev) / scale
        fields["residual"] = rhs * mask
    fields["matrix_scale"] = scale
    return Data(**fields)


def diagonal_of(edge_index: Array, block_values: Array, num_nodes: int) -> Array:
    """[N, b] scalar diagonal of a block edge list (what FolderDataset stores as ``diagonals``)."""
    if isinstance(block_values, torch.Tensor):
        ei = _t(edge_index, torch.long, block_values.device)
        bv = block_values.reshape(ei.shape[1], -1)
        b = int(round(bv.shape[1] ** 0.5))
        on = ei[0] == ei[1]
        out = torch.zeros(num_nodes, b, dtype=block_values.dtype, device=block_values.device)
        out[ei[0][on]] = bv[on].reshape(-1, b, b).diagonal(dim1=1, dim2=2)
        return out
    ei = np.asarray(edge_index)
    bv = np.asarray(block_values).reshape(ei.shape[1], -1)
    b = int(round(bv.shape[1] ** 0.5))
    on = ei[0] == ei[1]
    out = np.zeros((num_nodes, b), dtype=bv.dtype)
    out[ei[0][on]] = np.diagonal(bv[on].reshape(-1, b, b), axis1=1, axis2=2)
    return out
