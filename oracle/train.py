"""ORACLE (test infrastructure, not product): CPU restatement of the reference's training step.

Only tests/ may import this.  Plain torch autograd over the oracle forward (oracle/gnn.py), following

* SimpleTrainingWorkspace._shared_step   neural_cg/workspace.py:96-112   (G = gnn(x, ei, ea); d = AATPE(residual, G, mask))
* ScaledTrainingWorkspace._shared_step   neural_cg/scaled_workspace.py:96-112  (the same with diag = inv_diag)
* the losses of neural_cg/loss.py: rel_l2_loss :10-27, RelativeL2Loss_ANorm :168-188 (the config default,
  config/loss.yaml), L2Loss_ANorm :190-203, CosineSimilarityLoss_ANorm :228-241, ConjGradLoss_ANorm :280-308,
  NifLoss :119-139 -- each a per-sample mean over the ``ptr`` ranges of the batch.

Parity status: PINNED.  tests/golden/make_golden_train.py ran the UNMODIFIED reference workspaces + loss through torch
autograd in the build container; tests/test_oracle_golden.py checks this file against those losses and against the
gradient of every parameter (tests/golden/train_*.npz)."""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F

from . import gnn as ognn


def rel_l2(pred, gt, sqr_out=True, eps=1e-6):
    num, den = torch.linalg.norm((pred - gt).flatten()), torch.linalg.norm(gt.flatten())
    return num ** 2 / (den ** 2 + eps) if sqr_out else num / (den + eps)


def loss_value(name: str, batch, d: torch.Tensor, ptr) -> torch.Tensor:
    """neural_cg/loss.py forward of the named loss; ``batch`` has edge_index, matrix_values, mask, residual."""
    name = name.lower()
    ranges = list(zip(ptr[:-1], ptr[1:]))
    r = batch["residual"]
    if name == "nifloss_norm":
        Ar = ognn.graph_spmv(r, batch["edge_index"], batch["matrix_values"], None)
        return sum(rel_l2(d[b:e], Ar[b:e]) for b, e in ranges) / len(ranges)
    Ad = ognn.graph_spmv(d, batch["edge_index"], batch["matrix_values"], batch["mask"])
    if name == "relativel2loss_anorm":
        return sum(rel_l2(Ad[b:e], r[b:e]) for b, e in ranges) / len(ranges)
    if name in ("l2loss_anorm", "proploss", "relproploss"):
        return F.mse_loss(Ad, r)
    if name == "l1loss":
        return F.l1_loss(Ad, r)
    if name == "cosinesimilarityloss_anorm":
        return sum(1 - F.cosine_similarity(r[b:e].flatten(), Ad[b:e].flatten(), dim=-1, eps=1e-6) for b, e in ranges) / len(ranges)
    if name == "conjgradloss_anorm":
        def one(b, e):
            alpha = torch.dot(r[b:e].flatten(), d[b:e].flatten()) / (torch.dot(d[b:e].flatten(), Ad[b:e].flatten()) + 1e-6)
            return rel_l2(alpha * Ad[b:e], r[b:e])
        return sum(one(b, e) for b, e in ranges) / len(ranges)
    raise ValueError(f"Unknown loss {name}")


def training_step(sd: Dict[str, torch.Tensor], batch, epsilon: float, block_size: int, loss: str = "RelativeL2Loss_ANorm",
                  scaled: bool = False, prefix: str = "gnn."):
    """Returns (loss, {parameter name: gradient}, G).  ``sd`` holds plain tensors (copied; the caller's are untouched)."""
    params = {k: v.detach().clone().requires_grad_("node_msg_norm" not in k) for k, v in sd.items()}
    _, out = ognn.node_edge_processing(params, batch["x"], batch["edge_index"], batch["edge_attr"], None, prefix)
    G = out.reshape(-1, block_size, block_size)
    d = ognn.aatpe(batch["residual"], batch["edge_index"], G, epsilon, batch["mask"], batch["inv_diag"] if scaled else None)
    ptr = [int(v) for v in batch["ptr"]]
    val = loss_value(loss, batch, d, ptr)
    val.backward()
    return val.detach(), {k: p.grad for k, p in params.items() if p.requires_grad}, G.detach()
