"""Host-side mirror of ``neural_cg/nn/gnns.py``: ``NodeEdgeProcessing`` with the reference's constructor
signature (gnns.py:10-32) and ``state_dict`` key names, whose ``forward`` runs the
hand-written sm_100a kernels through the C ABI (``b200spai_gnn_forward``) instead of torch/PyG ops.

The parameters live in ordinary ``nn.Linear`` / ``nn.LayerNorm`` modules so that checkpoints of the
reference load unchanged (``load_state_dict``) and random initialisation consumes the torch RNG in the
same order (``self.apply(weight_init)``, gnns.py:75).  Those modules are never *called*: ``forward`` packs
their weights into one fp32 blob (``pack_state_dict``) and hands device pointers to the library.  There
is no eager fallback -- without a CUDA device ``forward`` raises.

Supported architecture = the reference's config/gnn.yaml: 16 features, ``num_layers: 2`` MLPs (3 Linear),
GELU, ``pre_norm: none`` on encoders/decoder, ``pre_norm: layer`` on the three MP MLPs, residuals on,
``aggr: add``; node/edge input widths <= 16, block size <= 4, any number of MP layers.  Anything else
raises ``NotImplementedError`` at construction.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Union

import numpy as np
import torch
from torch import nn

from ... import _lib
from ..utils.weight_init import weight_init
from .basic_layers import FeedForward, MPLayer

F = 16
MLP16 = 3 * (F * F + F)
MLP48 = 48 * F + F + 2 * (F * F + F)
LAYER_FLOATS = 2 * MLP48 + MLP16


def blob_length(num_layers: int) -> int:
    return 2 * MLP16 + num_layers * LAYER_FLOATS + MLP48


def _mlp_block(sd: Dict[str, torch.Tensor], prefix: str, k_pad: int, out_pad: int = F) -> np.ndarray:
    """One FeedForward as [W1^T (k_pad x 16), b1, W2^T, b2, W3^T (16 x 16, zero-padded outputs), b3] in fp32.
    A LayerNorm pre-norm (weight g, bias c) is folded into the first Linear in float64:
        W1'[k][j] = g[k] * W1[j][k],   b1'[j] = b1[j] + sum_k c[k] * W1[j][k]
    so the kernel only has to apply (v - mean) * rstd."""
    W1 = sd[prefix + "lift.0.weight"].detach().double().cpu().numpy()  # [16, in]
    b1 = sd[prefix + "lift.0.bias"].detach().double().cpu().numpy()
    if prefix + "pre_norm.weight" in sd:
        g = sd[prefix + "pre_norm.weight"].detach().double().cpu().numpy()
        c = sd[prefix + "pre_norm.bias"].detach().double().cpu().numpy()
        b1 = b1 + W1 @ c
        W1 = W1 * g[None, :]
    W2 = sd[prefix + "body.0.0.weight"].detach().double().cpu().numpy()
    b2 = sd[prefix + "body.0.0.bias"].detach().double().cpu().numpy()
    W3 = sd[prefix + "proj.0.weight"].detach().double().cpu().numpy()  # [out, 16]
    b3 = sd[prefix + "proj.0.bias"].detach().double().cpu().numpy()
    assert W1.shape[0] == F and W2.shape == (F, F) and W3.shape[1] == F and W1.shape[1] <= k_pad and W3.shape[0] <= out_pad
    W1t = np.zeros((k_pad, F))
    W1t[: W1.shape[1]] = W1.T
    W3t = np.zeros((F, out_pad))
    W3t[:, : W3.shape[0]] = W3.T
    b3p = np.zeros(out_pad)
    b3p[: b3.shape[0]] = b3
    return np.concatenate([W1t.ravel(), b1, W2.T.ravel(), b2, W3t.ravel(), b3p]).astype(np.float32)


def pack_state_dict(sd: Dict[str, torch.Tensor], num_layers: int, prefix: str = "") -> np.ndarray:
    """The "GNN weight blob" of include/b200spai.h (b200spai_gnn_create): node_enc, edge_enc,
    per layer [msg_mlp, edge_mlp, node_mlp], edge_dec."""
    parts = [_mlp_block(sd, prefix + "node_enc.", F), _mlp_block(sd, prefix + "edge_enc.", F)]
    for k in range(num_layers):
        p = f"{prefix}mp_layers.{k}."
        parts += [_mlp_block(sd, p + "msg_mlp.", 48), _mlp_block(sd, p + "edge_mlp.", 48), _mlp_block(sd, p + "node_mlp.", F)]
    parts.append(_mlp_block(sd, prefix + "edge_dec.", 48))
    blob = np.ascontiguousarray(np.concatenate(parts), dtype=np.float32)
    assert blob.size == blob_length(num_layers)
    return blob


def blob_layout(num_layers: int, prefix: str = ""):
    """(state_dict prefix, offset in the blob, padded input width) of every FeedForward in the weight blob."""
    out = [(prefix + "node_enc.", 0, F), (prefix + "edge_enc.", MLP16, F)]
    off = 2 * MLP16
    for k in range(num_layers):
        p = f"{prefix}mp_layers.{k}."
        out += [(p + "msg_mlp.", off, 48), (p + "edge_mlp.", off + MLP48, 48), (p + "node_mlp.", off + 2 * MLP48, F)]
        off += LAYER_FLOATS
    out.append((prefix + "edge_dec.", off, 48))
    return out


def pack_state_dict_torch(sd: Dict[str, torch.Tensor], num_layers: int, prefix: str = "") -> torch.Tensor:
    """``pack_state_dict`` with torch ops on the device the parameters live on (the blob after an optimiser step, no
    host round trip: ``b200spai_gnn_set_weights_dev``).  Same folding of the LayerNorm affine, in float64."""
    parts = []
    for pre, _, k_pad in blob_layout(num_layers, prefix):
        W1 = sd[pre + "lift.0.weight"].detach().double()
        b1 = sd[pre + "lift.0.bias"].detach().double()
        if pre + "pre_norm.weight" in sd:
            g, c = sd[pre + "pre_norm.weight"].detach().double(), sd[pre + "pre_norm.bias"].detach().double()
            b1 = b1 + W1 @ c
            W1 = W1 * g[None, :]
        W3 = sd[pre + "proj.0.weight"].detach().double()
        W1t = W1.new_zeros((k_pad, F))
        W1t[: W1.shape[1]] = W1.T
        W3t = W1.new_zeros((F, F))
        W3t[:, : W3.shape[0]] = W3.T
        b3 = W1.new_zeros(F)
        b3[: W3.shape[0]] = sd[pre + "proj.0.bias"].detach().double()
        parts += [W1t.reshape(-1), b1, sd[pre + "body.0.0.weight"].detach().double().T.reshape(-1),
                  sd[pre + "body.0.0.bias"].detach().double(), W3t.reshape(-1), b3]
    return torch.cat(parts).to(torch.float32).contiguous()


def unpack_grad_blob(gblob: torch.Tensor, sd: Dict[str, torch.Tensor], num_layers: int, prefix: str = "") -> Dict[str, torch.Tensor]:
    """dL/d(weight blob) (``b200spai_gnn_backward``) -> dL/d(state_dict parameter), the inverse of ``pack_state_dict``:
    transposes undone, padding dropped, and the LayerNorm affine (weight g, bias c) un-folded from the first Linear --
    with W1'[k][j] = g[k] W1[j][k] and b1'[j] = b1[j] + sum_k c[k] W1[j][k]:
        dW1[j][k] = g[k] dW1'[k][j] + db1'[j] c[k],   dg[k] = sum_j dW1'[k][j] W1[j][k],   dc[k] = sum_j db1'[j] W1[j][k]."""
    out: Dict[str, torch.Tensor] = {}
    for pre, off, k_pad in blob_layout(num_layers, prefix):
        W1 = sd[pre + "lift.0.weight"].detach()
        W3 = sd[pre + "proj.0.weight"].detach()
        n_in, n_out = W1.shape[1], W3.shape[0]
        o = off
        dW1p = gblob[o:o + k_pad * F].view(k_pad, F)[:n_in]
        o += k_pad * F
        db1p = gblob[o:o + F]
        o += F
        dW2t = gblob[o:o + F * F].view(F, F)
        o += F * F
        db2 = gblob[o:o + F]
        o += F
        dW3t = gblob[o:o + F * F].view(F, F)
        o += F * F
        db3 = gblob[o:o + F]
        if pre + "pre_norm.weight" in sd:
            g, c = sd[pre + "pre_norm.weight"].detach(), sd[pre + "pre_norm.bias"].detach()
            out[pre + "pre_norm.weight"] = (dW1p * W1.T).sum(dim=1)
            out[pre + "pre_norm.bias"] = W1.T @ db1p
            out[pre + "lift.0.weight"] = (dW1p * g[:, None]).T + db1p[:, None] * c[None, :]
        else:
            out[pre + "lift.0.weight"] = dW1p.T.clone()
        out[pre + "lift.0.bias"] = db1p.clone()
        out[pre + "body.0.0.weight"] = dW2t.T.clone()
        out[pre + "body.0.0.bias"] = db2.clone()
        out[pre + "proj.0.weight"] = dW3t[:, :n_out].T.clone()
        out[pre + "proj.0.bias"] = db3[:n_out].clone()
    return out


def _check_mlp(cfg: dict, pre_norm: str, what: str):
    ok = (str(cfg.get("pre_norm", "none")).lower() in pre_norm and int(cfg["hidden_channels"]) == F
          and int(cfg["num_layers"]) == 2 and cfg.get("activation", "gelu") == "gelu"
          and cfg.get("out_activation", "none") == "none")
    if not ok:
        raise NotImplementedError(f"{what}={dict(cfg)}: the CUDA path implements the reference default only "
                                  f"(hidden 16, num_layers 2, gelu, pre_norm in {pre_norm})")


def graph_key(edge_index: torch.Tensor, num_nodes: int):
    """Cache key of a topology: identity of the ``edge_index`` tensor's memory (address, layout, device) AND its in-place
    version counter.  Only valid while the cached graph keeps a reference to that tensor (``DeviceGraph._keepalive``):
    torch's caching allocator hands a freed block back at the same address, so an address alone says nothing about the
    connectivity stored there (two meshes of equal N and E routinely land on the same block)."""
    return (edge_index.data_ptr(), edge_index._version, edge_index.device.index, tuple(edge_index.stride()),
            int(num_nodes), int(edge_index.shape[1]))


class DeviceGraph:
    """Per-topology preprocessing (b200spai_graph_create): target-sorted edges, segment offsets, workspaces."""

    def __init__(self, edge_index: torch.Tensor, num_nodes: int):
        _lib.require_cuda()
        assert edge_index.is_cuda and edge_index.dtype == torch.int64 and edge_index.dim() == 2 and edge_index.shape[0] == 2
        self._lib = _lib.load()
        self.num_nodes, self.num_edges = int(num_nodes), int(edge_index.shape[1])
        ei = edge_index.contiguous()
        self._h = C.c_void_p()
        _lib.check(self._lib.b200spai_graph_create(self.num_nodes, self.num_edges, C.c_void_p(ei.data_ptr()),
                                                   C.c_void_p(torch.cuda.current_stream().cuda_stream), C.byref(self._h)),
                   "graph_create")
        self.key = graph_key(edge_index, num_nodes)
        self._keepalive = edge_index      # pins the storage the key names: it cannot be freed and recycled under the cache

    def close(self):
        if self._h:
            self._lib.b200spai_graph_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GnnTape:
    """What ONE training forward keeps for its backward (b200spai_gnn_tape_create): per-layer node features, edge
    features and aggregated messages; everything inside the MLPs is recomputed by the backward kernels."""

    def __init__(self, engine, graph: DeviceGraph):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        _lib.check(self._lib.b200spai_gnn_tape_create(engine, graph._h, C.byref(self._h)), "gnn_tape_create")

    def close(self):
        if self._h:
            self._lib.b200spai_gnn_tape_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class _GnnTrainFn(torch.autograd.Function):
    """NodeEdgeProcessing.forward under autograd (the reference gets this from torch autograd through PyG ops,
    gnns.py:77-97): forward = b200spai_gnn_forward_train (the inference kernels + a tape), backward =
    b200spai_gnn_backward (dL/dG -> dL/d weight blob) + ``unpack_grad_blob``.  The inputs x / edge_attr get no gradient
    (the reference never asks for one: they are data)."""

    @staticmethod
    def forward(ctx, module, graph, x, ea, names, *params):
        lib = _lib.load()
        engine = module._get_engine(device_pack=True)
        tape = GnnTape(engine, graph)
        E = graph.num_edges
        edge_out = torch.empty((E, module.edge_out), dtype=torch.float32, device=x.device)
        _lib.check(lib.b200spai_gnn_forward_train(engine, graph._h, tape._h, C.c_void_p(x.data_ptr()), C.c_void_p(ea.data_ptr()),
                                                  C.c_void_p(edge_out.data_ptr()),
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream)), "gnn_forward_train")
        ctx.module, ctx.graph, ctx.tape, ctx.x, ctx.ea, ctx.names = module, graph, tape, x, ea, names
        ctx.weights_version = module._weights_version()
        ctx.save_for_backward(*params)
        return edge_out

    @staticmethod
    def backward(ctx, grad_out):
        module = ctx.module
        if module._weights_version() != ctx.weights_version:
            raise _lib.B200SpaiError("the GNN weights changed between forward and backward (the backward recomputes the "
                                     "MLP internals from the weights: call backward() before optimizer.step())")
        lib = _lib.load()
        params = ctx.saved_tensors
        gG = grad_out.detach().to(torch.float32).contiguous()
        gblob = torch.empty(blob_length(module.num_mp_layers), dtype=torch.float32, device=gG.device)
        _lib.check(lib.b200spai_gnn_backward(module._engine, ctx.graph._h, ctx.tape._h, C.c_void_p(ctx.x.data_ptr()),
                                             C.c_void_p(ctx.ea.data_ptr()), C.c_void_p(gG.data_ptr()),
                                             C.c_void_p(gblob.data_ptr()),
                                             C.c_void_p(torch.cuda.current_stream().cuda_stream)), "gnn_backward")
        ctx.tape.close()
        grads = unpack_grad_blob(gblob, dict(zip(ctx.names, params)), module.num_mp_layers)
        return (None, None, None, None, None) + tuple(
            grads[n].to(p.dtype) if ctx.needs_input_grad[5 + i] else None for i, (n, p) in enumerate(zip(ctx.names, params)))


class NodeEdgeProcessing(nn.Module):
    def __init__(
        self,
        node_in_features: int,
        node_out_features: Union[int, None],
        node_encoder: dict,
        node_decoder: dict,
        edge_in_features: int,
        edge_out_features: int,
        edge_encoder: dict,
        edge_decoder: dict,
        num_mp_layers: int,
        node_features: int,
        edge_features: int,
        node_residual: bool,
        edge_residual: bool,
        node_mlp: dict,
        edge_mlp: dict,
        msg_mlp: dict,
        msg_norm: bool,
        aggr: str = "add",
    ):
        super().__init__()
        if node_out_features is not None:
            raise NotImplementedError("node decoder: the workspaces always pass node_out_features=None (workspace.py:72)")
        if not (node_features == F and edge_features == F and node_residual and edge_residual and aggr == "add"):
            raise NotImplementedError("the CUDA path implements 16 features, residual MP layers and aggr='add'")
        if not (1 <= node_in_features <= F and 1 <= edge_in_features <= F and 1 <= edge_out_features <= F):
            raise NotImplementedError("input/output widths must be in [1,16]")
        for cfg, pn, what in ((node_encoder, ("none",), "node_encoder"), (edge_encoder, ("none",), "edge_encoder"),
                              (edge_decoder, ("none",), "edge_decoder"),
                              (node_mlp, ("layer", "layernorm", "layer_norm"), "node_mlp"),
                              (edge_mlp, ("layer", "layernorm", "layer_norm"), "edge_mlp"),
                              (msg_mlp, ("layer", "layernorm", "layer_norm"), "msg_mlp")):
            _check_mlp(cfg, pn, what)
        # same registration order as the reference (gnns.py:34-74) -> same RNG consumption, same keys
        self.node_enc = FeedForward(in_channels=node_in_features, out_channels=node_features, **node_encoder)
        self.node_dec = nn.Identity()
        self.edge_enc = FeedForward(in_channels=edge_in_features, out_channels=edge_features, **edge_encoder)
        self.edge_dec = FeedForward(in_channels=edge_features + 2 * node_features, out_channels=edge_out_features, **edge_decoder)
        self.mp_layers = nn.ModuleList()
        for _ in range(num_mp_layers):
            self.mp_layers.append(MPLayer(node_channels=node_features, edge_channels=edge_features,
                                          node_residual=node_residual, edge_residual=edge_residual, node_mlp=node_mlp,
                                          edge_mlp=edge_mlp, msg_mlp=msg_mlp, aggr=aggr, msg_norm=msg_norm))
        self.apply(weight_init)
        self.node_in, self.edge_in, self.edge_out = node_in_features, edge_in_features, edge_out_features
        self.num_mp_layers = num_mp_layers
        self._engine = None
        self._engine_version = None
        self._graph: Optional[DeviceGraph] = None

    # ---- device engine ------------------------------------------------------------------------------
    def _weights_version(self):
        """Cheap fingerprint of the weights (storage pointer + in-place version counter of every parameter).  The
        parameter list is collected once -- walking the module tree costs ~0.7 ms per call, which is more than the whole
        GNN forward of a small mesh -- and again after ``_apply`` (``.to`` / ``.cuda`` / ``.float``).  Weights are
        expected to change through ``load_state_dict`` or in-place ops (both bump ``_version``)."""
        plist = self.__dict__.get("_plist")
        if plist is None:
            plist = list(self.parameters())
            self.__dict__["_plist"] = plist
        return tuple([(p.data_ptr(), p._version) for p in plist])

    def _apply(self, fn, *args, **kwargs):
        self.__dict__["_plist"] = None
        return super()._apply(fn, *args, **kwargs)

    def _get_engine(self, device_pack: bool = False):
        ver = self._weights_version()
        if device_pack and self._engine is not None and ver != self._engine_version:
            # training: the weights moved by an optimiser step -- repack on the device, no host round trip
            blob = pack_state_dict_torch(self.state_dict(), self.num_mp_layers)
            _lib.check(_lib.load().b200spai_gnn_set_weights_dev(self._engine, C.c_void_p(blob.data_ptr()), blob.numel(),
                                                                C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                       "gnn_set_weights_dev")
            self._engine_version = ver
            self._blob_keepalive = blob   # the copy is asynchronous on the current stream
            return self._engine
        if self._engine is None or ver != self._engine_version:
            self.release_engine()
            lib = _lib.load()
            blob = pack_state_dict(self.state_dict(), self.num_mp_layers)
            b = int(round(self.edge_out ** 0.5))
            if b * b != self.edge_out:
                raise NotImplementedError("edge_out_features must be block_size**2")
            h = C.c_void_p()
            _lib.check(lib.b200spai_gnn_create(self.node_in, self.edge_in, b, self.num_mp_layers,
                                               C.c_void_p(blob.ctypes.data), blob.size, C.byref(h)), "gnn_create")
            self._engine, self._engine_version = h, ver
        return self._engine

    def release_engine(self):
        if self._engine is not None:
            _lib.load().b200spai_gnn_destroy(self._engine)
            self._engine = None

    def graph_for(self, edge_index: torch.Tensor, num_nodes: int) -> DeviceGraph:
        """Topology preprocessing is cached on the identity of the edge_index tensor (fixed-topology datasets reuse one
        graph for every sample).  The cached graph holds that tensor alive and the key carries its version counter, so
        neither a recycled address nor an in-place edit can alias a stale topology (``graph_key``)."""
        key = graph_key(edge_index, num_nodes)
        if self._graph is None or self._graph.key != key:
            from . import basic_layers as _bl

            hit = _bl._graph_cache.get("g")
            self._graph = hit[1] if hit is not None and hit[0] == key else DeviceGraph(edge_index, num_nodes)
            # the edge-list operators of the same step (AATPE, the loss product) reuse this preprocessing
            _bl._graph_cache["g"] = (key, self._graph)
        return self._graph

    @property
    def launches(self) -> int:
        if self._engine is None:
            return 0
        v = C.c_int64()
        _lib.check(_lib.load().b200spai_gnn_launch_count(self._engine, C.byref(v)), "gnn_launch_count")
        return v.value

    def forward(self, node_attr: torch.Tensor, edge_index: torch.Tensor, edge_attr: torch.Tensor,
                graph: Optional[DeviceGraph] = None, autograd: bool = False):
        """``autograd=False`` (default): inference, no graph is recorded whatever the grad mode.  ``autograd=True`` (the
        workspaces' training / validation steps): the edge output carries a grad_fn whose backward runs the CUDA
        backward kernels; the node output is then ``None`` (the workspaces build the GNN with node_out_features=None and
        use the edge output only, workspace.py:92-94)."""
        _lib.require_cuda()
        if not (node_attr.is_cuda and edge_index.is_cuda and edge_attr.is_cuda):
            raise _lib.B200SpaiError("NodeEdgeProcessing.forward needs CUDA tensors: there is no CPU path")
        N, E = node_attr.shape[0], edge_index.shape[1]
        assert node_attr.shape[1] == self.node_in and edge_attr.shape == (E, self.edge_in)
        graph = graph or self.graph_for(edge_index, N)
        x = node_attr.detach().to(torch.float32).contiguous()
        ea = edge_attr.detach().to(torch.float32).contiguous()
        if autograd and torch.is_grad_enabled():
            named = [(n, p) for n, p in self.named_parameters() if "node_msg_norm" not in n]
            if any(p.requires_grad for _, p in named):
                names = tuple(n for n, _ in named)
                edge_out = _GnnTrainFn.apply(self, graph, x, ea, names, *[p for _, p in named])
                return None, edge_out
        return self._forward_inference(graph, x, ea, N, E)

    @torch.no_grad()
    def _forward_inference(self, graph, x, ea, N, E):
        edge_out = torch.empty((E, self.edge_out), dtype=torch.float32, device=x.device)
        node_out = torch.empty((N, F), dtype=torch.float32, device=x.device)
        lib = _lib.load()
        _lib.check(lib.b200spai_gnn_forward(self._get_engine(), graph._h, C.c_void_p(x.data_ptr()), C.c_void_p(ea.data_ptr()),
                                            C.c_void_p(edge_out.data_ptr()), C.c_void_p(node_out.data_ptr()),
                                            C.c_void_p(torch.cuda.current_stream().cuda_stream)), "gnn_forward")
        return node_out, edge_out

    def __del__(self):
        try:
            self.release_engine()
        except Exception:
            pass
