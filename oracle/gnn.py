"""ORACLE (test infrastructure, not product): CPU restatement of the reference GNN forward.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  It is a plain-torch fp32 restatement of

* FeedForward                neural_cg/nn/basic_layers.py:73-109  (pre_norm -> lift -> body* -> proj)
* get_activation / _normalization   basic_layers.py:13-44 (note "batch" maps to RMSNorm, :37-38)
* MPLayer                    basic_layers.py:145-225  (incl. the hasattr("msg_norm") quirk at :200,
                                                     which means MessageNorm is never applied)
* NodeEdgeProcessing         neural_cg/nn/gnns.py:77-97
* GraphSpmv / AATPE          basic_layers.py:112-142 / :228-261
* SimpleTrainingWorkspace.forward   neural_cg/workspace.py:92-94

with torch_geometric 2.6.1's MessagePassing conventions written out (flow="source_to_target":
x_j = x[edge_index[0]], x_i = x[edge_index[1]], sum at edge_index[1] with dim_size = N;
flow="target_to_source": swapped).  Aggregation is ``zeros.index_add_`` which on the CPU accumulates
sequentially in edge order, exactly like PyG's ``scatter_add_``.

Parameters are read from a ``state_dict`` with the reference's key names, so the
same weights drive the reference (when goldens are minted), this oracle and the CUDA kernels.

Parity status: PINNED against the reference's own basic_layers.py/gnns.py executed in the build
container with a torch_geometric stand-in (tests/golden/make_golden.py, fixtures in tests/golden/*.npz;
tests/test_oracle_golden.py).  torch_geometric itself is absent, so its MessagePassing glue is
restated on both sides; everything else on that side is the unmodified reference.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.nn.functional as F

DEFAULT_MLP = {"pre_norm": "none", "hidden_channels": 16, "num_layers": 2}
DEFAULT_MLP_LN = {"pre_norm": "layer", "hidden_channels": 16, "num_layers": 2}

DEFAULT_GNN_CFG = {  # config/gnn.yaml resolved
    "node_encoder": DEFAULT_MLP,
    "edge_encoder": DEFAULT_MLP,
    "node_decoder": DEFAULT_MLP,
    "edge_decoder": DEFAULT_MLP,
    "num_mp_layers": 4,
    "node_residual": True,
    "edge_residual": True,
    "node_features": 16,
    "edge_features": 16,
    "node_mlp": DEFAULT_MLP_LN,
    "edge_mlp": DEFAULT_MLP_LN,
    "msg_mlp": DEFAULT_MLP_LN,
    "msg_norm": True,
    "aggr": "add",
}


def _act(name: str, x: torch.Tensor) -> torch.Tensor:
    name = name.lower()
    if name == "gelu":
        return F.gelu(x)  # exact erf form, nn.GELU() default
    if name == "relu":
        return F.relu(x)
    if name == "tanh":
        return torch.tanh(x)
    if name == "sigmoid":
        return torch.sigmoid(x)
    if name == "elu":
        return F.elu(x)
    if name == "leaky_relu":
        return F.leaky_relu(x)
    if name == "none":
        return x
    raise ValueError(f"Activation {name} not supported.")


def _pre_norm(kind: str, sd: Dict[str, torch.Tensor], prefix: str, x: torch.Tensor) -> torch.Tensor:
    kind = kind.lower()
    if kind == "none":
        return x
    if kind in ("layer", "layernorm", "layer_norm"):
        return F.layer_norm(x, (x.shape[-1],), sd[prefix + "pre_norm.weight"], sd[prefix + "pre_norm.bias"], 1e-5)
    if kind in ("batch", "batchnorm", "batch_norm", "rms", "rmsnorm", "rms_norm"):
        return F.rms_norm(x, (x.shape[-1],), sd[prefix + "pre_norm.weight"], None)
    raise ValueError(f"Normalization {kind} not supported.")


def feed_forward(sd, prefix, x, cfg, activation="gelu", out_activation="none"):
    """basic_layers.py:103-109."""
    x = _pre_norm(cfg.get("pre_norm", "none"), sd, prefix, x)
    x = _act(activation, F.linear(x, sd[prefix + "lift.0.weight"], sd[prefix + "lift.0.bias"]))
    for k in range(cfg["num_layers"] - 1):
        x = _act(activation, F.linear(x, sd[f"{prefix}body.{k}.0.weight"], sd[f"{prefix}body.{k}.0.bias"]))
    return _act(out_activation, F.linear(x, sd[prefix + "proj.0.weight"], sd[prefix + "proj.0.bias"]))


def _scatter_sum(src: torch.Tensor, index: torch.Tensor, n: int) -> torch.Tensor:
    return src.new_zeros((n,) + tuple(src.shape[1:])).index_add_(0, index, src)


def mp_layer(sd, prefix, node, edge_index, edge, cfg):
    """basic_layers.py:193-225; returns (node_out, edge_out)."""
    src, dst = edge_index[0], edge_index[1]
    x_i, x_j = node[dst], node[src]
    feat = torch.cat([x_i, x_j, edge], dim=-1)
    msg = feed_forward(sd, prefix + "msg_mlp.", feat, cfg["msg_mlp"])
    agg = _scatter_sum(msg, dst, node.shape[0])
    if cfg.get("aggr", "add") == "mean":
        cnt = _scatter_sum(torch.ones_like(dst, dtype=node.dtype), dst, node.shape[0]).clamp(min=1)
        agg = agg / cnt[:, None]
    node_new = feed_forward(sd, prefix + "node_mlp.", agg, cfg["node_mlp"])
    # basic_layers.py:200 tests hasattr(self, "msg_norm"); the attribute is node_msg_norm -> never taken
    node_out = node + node_new if cfg["node_residual"] else node_new
    edge_new = feed_forward(sd, prefix + "edge_mlp.", feat, cfg["edge_mlp"])  # uses the OLD node features
    edge_out = edge + edge_new if cfg["edge_residual"] else edge_new
    return node_out, edge_out


def node_edge_processing(sd, x, edge_index, edge_attr, cfg=None, prefix="gnn."):
    """gnns.py:77-97 with node_out_features=None; returns (node_out [N,F], edge_out [E,b*b])."""
    cfg = cfg or DEFAULT_GNN_CFG
    node = feed_forward(sd, prefix + "node_enc.", x, cfg["node_encoder"])
    edge = feed_forward(sd, prefix + "edge_enc.", edge_attr, cfg["edge_encoder"])
    for k in range(cfg["num_mp_layers"]):
        node, edge = mp_layer(sd, f"{prefix}mp_layers.{k}.", node, edge_index, edge, cfg)
    dec_in = torch.cat([edge, node[edge_index[0]], node[edge_index[1]]], dim=-1)
    return node, feed_forward(sd, prefix + "edge_dec.", dec_in, cfg["edge_decoder"])


def workspace_forward(sd, x, edge_index, edge_attr, block_size, cfg=None, prefix="gnn."):
    """workspace.py:92-94 -> [E,b,b]."""
    with torch.no_grad():
        _, out = node_edge_processing(sd, x, edge_index, edge_attr, cfg, prefix)
    return out.reshape(-1, block_size, block_size)


def workspace_forward_chunked(sd, x, edge_index, edge_attr, block_size, cfg=None, prefix="gnn.", chunk=1 << 22):
    """Same arithmetic as ``workspace_forward`` with the per-edge work done ``chunk`` edges at a time, so that the
    [E,48] concatenations (29 GB each at 216^3) never exist.  Every per-edge row goes through the same torch CPU
    kernels and messages are added into the aggregate in the same ascending edge order (sequential ``index_add_``),
    so the result equals the un-chunked forward up to torch's CPU GELU using its scalar erf for the last few elements
    of every thread's range (where those fall depends on the tensor length): a handful of entries move by one fp32
    ulp of an intermediate, |dG| <= 1e-7 (measured by tests/golden/make_golden_c5.py and tests/test_oracle_golden.py)."""
    cfg = cfg or DEFAULT_GNN_CFG
    E = edge_index.shape[1]
    src, dst = edge_index[0], edge_index[1]
    with torch.no_grad():
        node = feed_forward(sd, prefix + "node_enc.", x, cfg["node_encoder"])
        edge = torch.empty((E, cfg["edge_features"]), dtype=x.dtype)
        for s in range(0, E, chunk):
            edge[s:s + chunk] = feed_forward(sd, prefix + "edge_enc.", edge_attr[s:s + chunk], cfg["edge_encoder"])
        for k in range(cfg["num_mp_layers"]):
            p = f"{prefix}mp_layers.{k}."
            agg = node.new_zeros(node.shape)
            for s in range(0, E, chunk):
                sl = slice(s, min(s + chunk, E))
                feat = torch.cat([node[dst[sl]], node[src[sl]], edge[sl]], dim=-1)
                agg.index_add_(0, dst[sl], feed_forward(sd, p + "msg_mlp.", feat, cfg["msg_mlp"]))
                e_new = feed_forward(sd, p + "edge_mlp.", feat, cfg["edge_mlp"])   # OLD node features
                edge[sl] = edge[sl] + e_new if cfg["edge_residual"] else e_new
            assert cfg.get("aggr", "add") == "add"
            node_new = feed_forward(sd, p + "node_mlp.", agg, cfg["node_mlp"])
            node = node + node_new if cfg["node_residual"] else node_new
        out = torch.empty((E, block_size * block_size), dtype=x.dtype)
        for s in range(0, E, chunk):
            sl = slice(s, min(s + chunk, E))
            dec_in = torch.cat([edge[sl], node[src[sl]], node[dst[sl]]], dim=-1)
            out[sl] = feed_forward(sd, prefix + "edge_dec.", dec_in, cfg["edge_decoder"])
    return out.reshape(-1, block_size, block_size)


def graph_spmv(X, edge_index, A, mask=None, transpose=False):
    """basic_layers.py:112-142.  X [N,b], A [E,b,b]."""
    row, col = edge_index[0], edge_index[1]
    if transpose:  # flow source_to_target: x_j = X[row], summed at col
        out = _scatter_sum(torch.bmm(A.transpose(-1, -2), X[row].unsqueeze(-1)).squeeze(-1), col, X.shape[0])
    else:          # flow target_to_source: x_j = X[col], summed at row
        out = _scatter_sum(torch.bmm(A, X[col].unsqueeze(-1)).squeeze(-1), row, X.shape[0])
    return out * mask if mask is not None else out


def aatpe(x, edge_index, G, epsilon, mask=None, diag: Optional[torch.Tensor] = None):
    """basic_layers.py:235-261: eps*x + G (G^T x), masked; optionally Jacobi-scaled."""
    t = graph_spmv(x, edge_index, G, mask, transpose=True)
    eps_x = epsilon * x
    if diag is not None:
        t = t * diag
        eps_x = eps_x * diag
    return eps_x + graph_spmv(t, edge_index, G, mask)


def state_dict_keys(cfg=None, prefix="gnn."):
    """The parameter names the reference workspace exposes."""
    cfg = cfg or DEFAULT_GNN_CFG
    keys = []

    def ff(p, c):
        if c.get("pre_norm", "none") != "none":
            keys.append(p + "pre_norm.weight")
            if c["pre_norm"].startswith("layer"):
                keys.append(p + "pre_norm.bias")
        keys.extend([p + "lift.0.weight", p + "lift.0.bias"])
        for k in range(c["num_layers"] - 1):
            keys.extend([f"{p}body.{k}.0.weight", f"{p}body.{k}.0.bias"])
        keys.extend([p + "proj.0.weight", p + "proj.0.bias"])

    ff(prefix + "node_enc.", cfg["node_encoder"])
    ff(prefix + "edge_enc.", cfg["edge_encoder"])
    ff(prefix + "edge_dec.", cfg["edge_decoder"])
    for k in range(cfg["num_mp_layers"]):
        for m in ("node_mlp", "edge_mlp", "msg_mlp"):
            ff(f"{prefix}mp_layers.{k}.{m}.", cfg[m])
        if cfg.get("msg_norm", True):
            keys.append(f"{prefix}mp_layers.{k}.node_msg_norm.scale")
    return keys
