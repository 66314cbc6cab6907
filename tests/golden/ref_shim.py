"""Stand-in modules that let the UNMODIFIED reference sources under /root/reference be imported
in this container, where torch_geometric / lightning / hydra are not installed.

Only used by ``make_golden.py`` (fixture generation, run once in the build container).  Nothing in the
product, the tests or the bench imports this file at run time: /root/reference does not exist on the
GPU box, and the fixtures it produced are committed under tests/golden/.

What is real and what is shimmed when the goldens are generated:

* REAL (executed from /root/reference as-is): neural_cg/nn/basic_layers.py (FeedForward, MPLayer,
  GraphSpmv, AATPE), neural_cg/nn/gnns.py (NodeEdgeProcessing), neural_cg/data.py (make_data,
  make_bsr_from_coo_inds, apply_dbc_masking), neural_cg/utils/validate.py (to_csr_cpu,
  get_pcg_iter_time_scipy, get_cg_iter_time_scipy), datagen/synthetic.py
  (generate_spd_sparse_matrix), torch 2.11 CPU kernels, scipy 1.18.1.
* SHIMMED (restated here, torch_geometric 2.6.1 is absent): ``MessagePassing.propagate`` /
  ``edge_updater`` argument collection + sum aggregation, ``MessageNorm``, ``utils.scatter``,
  ``data.Data``.  The conventions restated are the documented PyG ones: with
  ``flow="source_to_target"`` ``x_j = x[edge_index[0]]``, ``x_i = x[edge_index[1]]`` and messages are
  summed at ``edge_index[1]``; ``flow="target_to_source"`` swaps the two; ``aggr="add"`` is
  ``zeros(dim_size, F).scatter_add_(0, index, src)``; ``update`` receives the aggregate first.
"""
from __future__ import annotations

import inspect
import sys
import types

import torch
from torch import nn


class MessagePassing(nn.Module):
    def __init__(self, aggr="add", flow="source_to_target", node_dim=-2, **kwargs):
        super().__init__()
        assert aggr in ("add", "sum", "mean")
        self.aggr = aggr
        self.flow = flow
        self.node_dim = node_dim

    def reset_parameters(self):  # PyG resets its aggregation module here; no RNG use
        pass

    def _ij(self):
        return (1, 0) if self.flow == "source_to_target" else (0, 1)

    def _collect(self, fn, edge_index, kwargs, skip=0):
        i, j = self._ij()
        out = {}
        names = list(inspect.signature(fn).parameters)[skip:]
        for name in names:
            if name.endswith("_i") or name.endswith("_j"):
                base = kwargs[name[:-2]]
                idx = edge_index[i] if name.endswith("_i") else edge_index[j]
                out[name] = base.index_select(0, idx)
            elif name in kwargs:
                out[name] = kwargs[name]
        return out

    def propagate(self, edge_index, size=None, **kwargs):
        i, _ = self._ij()
        msg = self.message(**self._collect(self.message, edge_index, kwargs))
        x = kwargs.get("x")
        dim_size = x.size(0)
        index = edge_index[i]
        out = msg.new_zeros((dim_size,) + tuple(msg.shape[1:]))
        out.scatter_add_(0, index.view(-1, *([1] * (msg.dim() - 1))).expand_as(msg), msg)
        if self.aggr == "mean":
            cnt = msg.new_zeros(dim_size).scatter_add_(0, index, msg.new_ones(index.numel()))
            out = out / cnt.clamp(min=1).view(-1, *([1] * (msg.dim() - 1)))
        upd_kwargs = self._collect(self.update, edge_index, kwargs, skip=1)
        return self.update(out, **upd_kwargs)

    def edge_updater(self, edge_index, size=None, **kwargs):
        return self.edge_update(**self._collect(self.edge_update, edge_index, kwargs))

    def message(self, x_j):
        return x_j

    def update(self, aggr_out):
        return aggr_out


class MessageNorm(nn.Module):
    def __init__(self, learn_scale: bool = False):
        super().__init__()
        self.scale = nn.Parameter(torch.empty(1), requires_grad=learn_scale)
        self.reset_parameters()

    def reset_parameters(self):
        self.scale.data.fill_(1.0)

    def forward(self, x, msg, p=2.0):
        msg = torch.nn.functional.normalize(msg, p=p, dim=-1)
        return msg * x.norm(p=p, dim=-1, keepdim=True) * self.scale


def scatter(src, index, dim=0, dim_size=None, reduce="sum"):
    assert dim == 0
    shape = (dim_size,) + tuple(src.shape[1:])
    idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    if reduce in ("sum", "add"):
        return src.new_zeros(shape).scatter_add_(0, idx, src)
    if reduce == "mean":
        cnt = src.new_zeros(dim_size).scatter_add_(0, index, src.new_ones(index.numel()))
        out = src.new_zeros(shape).scatter_add_(0, idx, src)
        return out / cnt.clamp(min=1).view(-1, *([1] * (src.dim() - 1)))
    raise NotImplementedError(reduce)


class Data:
    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    def to_dict(self):
        return dict(self.__dict__)

    def to(self, device):
        for k, v in self.__dict__.items():
            if isinstance(v, torch.Tensor):
                setattr(self, k, v.to(device))
        return self


class Dataset:
    def __init__(self, *a, **k):
        pass

    def __len__(self):
        return self.len()


def _unavailable(*a, **k):
    raise NotImplementedError("not shimmed")


def install(reference_root="/root/reference"):
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    tg = mod("torch_geometric")
    tg.nn = mod("torch_geometric.nn", MessagePassing=MessagePassing, MessageNorm=MessageNorm)
    tg.utils = mod(
        "torch_geometric.utils",
        scatter=scatter,
        coalesce=_unavailable,
        remove_self_loops=_unavailable,
        to_torch_coo_tensor=_unavailable,
        to_edge_index=_unavailable,
    )
    tg.data = mod("torch_geometric.data", Data=Data, Dataset=Dataset)
    tg.loader = mod("torch_geometric.loader", DataLoader=_unavailable)

    class _LModule(nn.Module):
        def save_hyperparameters(self, *a, **k):
            pass

        def log_dict(self, *a, **k):
            pass

        def log(self, *a, **k):
            pass

    mod("lightning", LightningModule=_LModule, LightningDataModule=object)

    def _hydra_main(*a, **k):
        return lambda fn: fn

    mod("hydra", main=_hydra_main)
    if reference_root not in sys.path:
        sys.path.insert(0, reference_root)
