"""ORACLE (test infrastructure, not product): CPU restatement of the reference's training step.

Only tests/ may import this.  Plain torch autograd over the oracle forward (oracle/gnn.py), following

* SimpleTrainingWorkspace._shared_step   neural_cg/workspace.py:96-112   (G = gnn(x, ei, ea); d = AATPE(residual, G, mask))
* ScaledTrainingWorkspace._shared_step   neural_cg/scaled_workspace.py:96-112  (the same with diag = inv_diag)
* every loss of neural_cg/loss.py that ``create_loss_item`` (:420-459) can return: rel_l2_loss :10-27 and the twelve names
  (RelativeL2Loss_ANorm :168-188 is the config default, config/loss.yaml) -- per-sample means over the ``ptr`` ranges of
  the batch, or one range with ``batch_less``.

Parity status: PINNED.  tests/golden/make_golden_train.py ran the UNMODIFIED reference workspaces + loss through torch
autograd in the build container; tests/test_oracle_golden.py checks this file against those losses and against the
gradient of every parameter (tests/golden/train_*.npz), and every loss name against the value and dL/dd the reference's
own loss classes give (tests/golden/train_losses.npz, minted by tests/golden/make_golden_losses.py)."""
from __future__ import annotations

from typing import Dict

import torch
import torch.nn.functional as F

from . import gnn as ognn


def rel_l2(pred, gt, sqr_out=True, eps=1e-6):
    num, den = torch.linalg.norm((pred - gt).flatten()), torch.linalg.norm(gt.flatten())
    return num ** 2 / (den ** 2 + eps) if sqr_out else num / (den + eps)


def loss_value(name: str, batch, d: torch.Tensor, ptr, batch_less: bool = False) -> torch.Tensor:
    """neural_cg/loss.py forward of the named loss (every name create_loss_item accepts, loss.py:420-459); ``batch`` has
    edge_index, matrix_values, mask, residual (and gt for the PlainNorm losses).  ``batch_less``: one range over the whole
    batch instead of the ``ptr`` ranges (loss.py:176-180)."""
    name = name.lower()
    ranges = [(0, int(d.shape[0]))] if batch_less else list(zip(ptr[:-1], ptr[1:]))
    r = batch["residual"]

    def mean(fn):
        return sum(fn(b, e) for b, e in ranges) / len(ranges)

    def cos(pred, gt):
        return 1 - F.cosine_similarity(gt.flatten(), pred.flatten(), dim=-1, eps=1e-6)

    if name == "relativel2loss_plainnorm":                                                   # loss.py:142-165
        return mean(lambda b, e: rel_l2(d[b:e], batch["gt"][b:e]))
    if name == "cosinesimilarityloss_plainnorm":                                             # loss.py:205-225
        return mean(lambda b, e: cos(d[b:e], batch["gt"][b:e]))
    if name == "nifloss_norm":                                                               # loss.py:119-139 (unmasked A r)
        Ar = ognn.graph_spmv(r, batch["edge_index"], batch["matrix_values"], None)
        return mean(lambda b, e: rel_l2(d[b:e], Ar[b:e]))
    Ad = ognn.graph_spmv(d, batch["edge_index"], batch["matrix_values"], batch["mask"])
    if name == "relativel2loss_anorm":                                                       # loss.py:168-188
        return mean(lambda b, e: rel_l2(Ad[b:e], r[b:e]))
    if name in ("l2loss_anorm", "proploss", "relproploss"):                                  # loss.py:190-203, 343-402
        return F.mse_loss(Ad, r)
    if name == "l1loss":                                                                     # loss.py:404-417
        return F.l1_loss(Ad, r)
    if name == "cosinesimilarityloss_anorm":                                                 # loss.py:228-241
        return mean(lambda b, e: cos(Ad[b:e], r[b:e]))

    def alpha(b, e):                                                                         # cg_alpha, loss.py:72-87
        return torch.dot(r[b:e].flatten(), d[b:e].flatten()) / (torch.dot(d[b:e].flatten(), Ad[b:e].flatten()) + 1e-6)

    if name == "conjgradloss_plainnorm":                                                     # loss.py:244-277
        return mean(lambda b, e: rel_l2(alpha(b, e) * d[b:e], batch["gt"][b:e]))
    if name == "conjgradloss_anorm":                                                         # loss.py:280-308
        return mean(lambda b, e: rel_l2(alpha(b, e) * Ad[b:e], r[b:e]))
    if name == "conjgradloss_anorm_norelative":                                              # loss.py:310-340
        return mean(lambda b, e: F.mse_loss(alpha(b, e) * Ad[b:e], r[b:e]))
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
