"""Host-side mirror of ``neural_cg/nn/basic_layers.py``.

``FeedForward`` (:73-109) and ``MPLayer`` (:145-225) are PARAMETER CONTAINERS here: they register the
same sub-modules in the same order as the reference, so ``state_dict`` keys, checkpoint loading and the
RNG consumption of random initialisation are identical, but their arithmetic is executed by the fused
kernels behind ``NodeEdgeProcessing.forward`` (gnns.py) -- calling them directly raises, because this
package has no eager PyTorch path.

``GraphSpmv`` (:112-142) and ``AATPE`` (:228-261) are the fp32 edge-list operators that define
M^-1 = G G^T + eps I at tensor level; they run ``b200spai_graph_spmv_f32`` (deterministic segmented
sums in the reference's CPU accumulation order) and keep the reference call signatures.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Union

import torch
from torch import nn

from ... import _lib

_ACTIVATIONS = {"relu": nn.ReLU, "tanh": nn.Tanh, "sigmoid": nn.Sigmoid, "gelu": nn.GELU, "elu": nn.ELU,
                "leaky_relu": nn.LeakyReLU, "none": nn.Identity}


def get_activation(activation: str) -> nn.Module:
    try:
        return _ACTIVATIONS[activation.lower()]()
    except KeyError:
        raise ValueError(f"Activation {activation} not supported.") from None


def get_normalization(normalization: str, channels: int) -> nn.Module:
    kind = normalization.lower()
    if kind == "none":
        return nn.Identity()
    if kind in ("layer", "layernorm", "layer_norm"):
        return nn.LayerNorm(channels)
    # the reference maps "batch*" to RMSNorm as well (basic_layers.py:37-38)
    if kind in ("batch", "batchnorm", "batch_norm", "rms", "rmsnorm", "rms_norm"):
        return nn.RMSNorm(channels)
    raise ValueError(f"Normalization {normalization} not supported.")


class _NoEagerPath(nn.Module):
    def forward(self, *args, **kwargs):
        raise _lib.B200SpaiError(
            f"{type(self).__name__} is a parameter container: its arithmetic runs inside the fused CUDA kernels of "
            "NodeEdgeProcessing.forward (libb200spai.so); there is no eager PyTorch path")


class FeedForward(_NoEagerPath):
    def __init__(self, in_channels: int, out_channels: int, hidden_channels: int, num_layers: int,
                 pre_norm: str = "none", activation: str = "gelu", out_activation: str = "none"):
        super().__init__()
        self.pre_norm = get_normalization(pre_norm, in_channels)
        self.lift = nn.Sequential(nn.Linear(in_channels, hidden_channels), get_activation(activation))
        self.body = nn.ModuleList(
            nn.Sequential(nn.Linear(hidden_channels, hidden_channels), get_activation(activation))
            for _ in range(1, num_layers))
        self.proj = nn.Sequential(nn.Linear(hidden_channels, out_channels), get_activation(out_activation))


class MessageNorm(nn.Module):
    """Stand-in for torch_geometric.nn.MessageNorm(learn_scale=False): owns the frozen ``scale`` parameter
    that appears in reference checkpoints.  The reference never applies it (basic_layers.py:200 tests
    ``hasattr(self, "msg_norm")`` while the attribute is ``node_msg_norm``), so neither do the kernels."""

    def __init__(self, learn_scale: bool = False):
        super().__init__()
        self.scale = nn.Parameter(torch.empty(1), requires_grad=learn_scale)
        self.reset_parameters()

    def reset_parameters(self):
        self.scale.data.fill_(1.0)


class MPLayer(_NoEagerPath):
    def __init__(self, node_channels: int, edge_channels: int, node_residual: bool, edge_residual: bool,
                 node_mlp: dict, edge_mlp: dict, msg_mlp: dict, aggr: str = "add", msg_norm: bool = True):
        super().__init__()
        self.aggr = aggr
        self.node_mlp = FeedForward(in_channels=node_channels, out_channels=node_channels, **node_mlp)
        self.edge_mlp = FeedForward(in_channels=2 * node_channels + edge_channels, out_channels=edge_channels, **edge_mlp)
        self.msg_mlp = FeedForward(in_channels=edge_channels + 2 * node_channels, out_channels=node_channels, **msg_mlp)
        self.node_residual = node_residual
        self.edge_residual = edge_residual
        if msg_norm:
            self.node_msg_norm = MessageNorm()


# ---- edge-list operators -----------------------------------------------------------------------------
_graph_cache = {}


def _graph_for(edge_index: torch.Tensor, num_nodes: int):
    """DeviceGraph cached on the identity of the edge_index tensor (one entry: the operators are called repeatedly on
    the same batch).  The entry keeps the tensor alive and is keyed on address + version + device (``gnns.graph_key``)."""
    from .gnns import DeviceGraph, graph_key

    key = graph_key(edge_index, num_nodes)
    hit = _graph_cache.get("g")
    if hit is None or hit[0] != key:
        _graph_cache["g"] = (key, DeviceGraph(edge_index, num_nodes))
    return _graph_cache["g"][1]


def _spmv_raw(graph, b: int, a: torch.Tensor, x: torch.Tensor, m: Optional[torch.Tensor], transpose: bool) -> torch.Tensor:
    """b200spai_graph_spmv_f32 on contiguous fp32 device tensors: y[N,b] = m * (A x)  (A^T x when ``transpose``)."""
    y = torch.empty((graph.num_nodes, b), dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().b200spai_graph_spmv_f32(graph._h, b, C.c_void_p(a.data_ptr()), C.c_void_p(x.data_ptr()),
                                                   C.c_void_p(m.data_ptr() if m is not None else 0), int(transpose),
                                                   C.c_void_p(y.data_ptr()),
                                                   C.c_void_p(torch.cuda.current_stream().cuda_stream)), "graph_spmv_f32")
    return y


class _GraphSpmvFn(torch.autograd.Function):
    """GraphSpmv under autograd (the reference differentiates PyG's propagate + bmm, basic_layers.py:130-142).  With
    gy' = mask * gy:  dL/dx = the operator with the OTHER transpose flag applied to gy' (b200spai_graph_spmv_f32),
    dL/dA_e = gy'[row_e] (x) x[col_e]  (transpose: x[row_e] (x) gy'[col_e])  (b200spai_graph_outer_f32)."""

    @staticmethod
    def forward(ctx, graph, b, transpose, x, a, m):
        ctx.graph, ctx.b, ctx.transpose, ctx.has_mask = graph, b, transpose, m is not None
        ctx.save_for_backward(x, a, m if m is not None else x.new_empty(0))
        return _spmv_raw(graph, b, a, x, m, transpose)

    @staticmethod
    def backward(ctx, gy):
        x, a, m = ctx.saved_tensors
        gy = gy.detach().to(torch.float32).reshape(-1, ctx.b)
        gy = (gy * m).contiguous() if ctx.has_mask else gy.contiguous()
        gx = ga = None
        if ctx.needs_input_grad[3]:
            gx = _spmv_raw(ctx.graph, ctx.b, a, gy, None, not ctx.transpose)
        if ctx.needs_input_grad[4]:
            ga = torch.empty_like(a)
            _lib.check(_lib.load().b200spai_graph_outer_f32(ctx.graph._h, ctx.b, C.c_void_p(gy.data_ptr()), C.c_void_p(x.data_ptr()),
                                                            int(ctx.transpose), 0, C.c_void_p(ga.data_ptr()),
                                                            C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                       "graph_outer_f32")
        return None, None, None, gx, ga, None


class GraphSpmv(nn.Module):
    """y_i = sum_j A_ij x_j on an edge list (``use_transpose``: y_j = sum_i A_ij^T x_i), then ``* mask``.
    Differentiable w.r.t. X and A when either requires grad (training); plain kernel call otherwise."""

    def __init__(self, use_transpose: bool = False):
        super().__init__()
        self.transpose = use_transpose

    def forward(self, X: torch.Tensor, edge_index: torch.Tensor, A: torch.Tensor, mask: Optional[torch.Tensor] = None,
                graph=None) -> torch.Tensor:
        _lib.require_cuda()
        if not (X.is_cuda and edge_index.is_cuda and A.is_cuda):
            raise _lib.B200SpaiError("GraphSpmv needs CUDA tensors: there is no CPU path")
        N = X.shape[0]
        b = 1 if A.dim() == 1 else A.shape[-1]
        assert X.numel() == N * b and A.numel() == edge_index.shape[1] * b * b
        graph = graph or _graph_for(edge_index, N)
        x = X.to(torch.float32).reshape(N, b).contiguous()
        a = A.to(torch.float32).contiguous()
        m = None if mask is None else mask.detach().to(torch.float32).expand(N, b).contiguous()
        if torch.is_grad_enabled() and (x.requires_grad or a.requires_grad):
            y = _GraphSpmvFn.apply(graph, b, bool(self.transpose), x, a, m)
        else:
            y = _spmv_raw(graph, b, a.detach(), x.detach(), m, bool(self.transpose))
        return y.reshape(X.shape)


class AATPE(nn.Module):
    """y = eps x + G (G^T x)  (with ``diag``: eps diag x + G (diag (G^T x)))  -- basic_layers.py:235-261.
    Differentiable w.r.t. ``boo_values`` (through GraphSpmv) when they require grad."""

    def __init__(self, epsilon):
        super().__init__()
        self.spmv = GraphSpmv()
        self.spmv_t = GraphSpmv(use_transpose=True)
        self.epsilon = epsilon

    def forward(self, x: torch.Tensor, edge_index, boo_values, mask: Union[torch.Tensor, None] = None,
                diag: Optional[torch.Tensor] = None):
        graph = _graph_for(edge_index, x.shape[0])
        AT_x = self.spmv_t(X=x, edge_index=edge_index, A=boo_values, mask=mask, graph=graph)
        eps_x = self.epsilon * x
        if diag is not None:
            assert diag.shape == AT_x.shape
            AT_x = AT_x * diag
            eps_x = eps_x * diag
        return eps_x + self.spmv(X=AT_x, edge_index=edge_index, A=boo_values, mask=mask, graph=graph)


class LLT(nn.Module):
    """y = L (L^T x)  (basic_layers.py:264-275)."""

    def __init__(self):
        super().__init__()
        self.spmv = GraphSpmv()
        self.spmv_t = GraphSpmv(use_transpose=True)

    def forward(self, x, edge_index, boo_values, mask: Union[torch.Tensor, None] = None):
        return self.spmv(X=self.spmv_t(X=x, edge_index=edge_index, A=boo_values, mask=mask), edge_index=edge_index,
                         A=boo_values, mask=mask)


class ToLowerTriangular(nn.Module):
    """Keep the edges with col <= row (basic_layers.py:278-294): the ``pattern="tril"`` switch."""

    def __init__(self, inplace: bool = False):
        super().__init__()
        self.inplace = inplace

    def forward(self, data, edge_index):
        keep = edge_index[1] <= edge_index[0]
        return data[keep], edge_index[:, keep]
