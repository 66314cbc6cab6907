"""Host-side mirror of ``neural_cg/loss.py``: the training losses on d = M^-1 r (``create_loss_item`` and the
classes it returns, same names and constructor keywords), for training through the fused kernels.

Every loss is a few reductions over [N, b] vectors (torch ops on device tensors: plumbing) around ONE edge-list
product ``A d`` -- ``GraphSpmv`` (nn/basic_layers.py), which runs ``b200spai_graph_spmv_f32`` and is differentiable
through ``b200spai_graph_spmv_f32`` / ``b200spai_graph_outer_f32``.  The default of ``config/loss.yaml`` is
``RelativeL2Loss_ANorm``: mean over the samples of the batch of ||A d - r||^2 / (||r||^2 + eps)  (loss.py:168-188).

A batch is split into samples by ``batch.ptr`` unless ``batch_less`` (loss.py:176-180).  The reference's quirks are kept:
``PropLoss`` / ``RelPropLoss`` compute a per-sample quantity and then return the plain MSE (loss.py:343-402).
"""
from __future__ import annotations

from typing import Callable, List

import torch
from torch import nn
from torch.nn import functional as F

from .nn.basic_layers import GraphSpmv


def _check(pred: torch.Tensor, gt: torch.Tensor) -> None:
    if gt.shape != pred.shape:
        raise ValueError(f"Shape mismatch: {gt.shape} != {pred.shape}")


def rel_l2_loss(pred, gt, sqr_out=True, eps=1e-6):
    """||pred - gt||^2 / (||gt||^2 + eps)  (``sqr_out=False``: the ratio of the norms), loss.py:10-27."""
    _check(pred, gt)
    num, den = torch.linalg.norm((pred - gt).flatten()), torch.linalg.norm(gt.flatten())
    return num ** 2 / (den ** 2 + eps) if sqr_out else num / (den + eps)


nif_loss = rel_l2_loss  # loss.py:43-60: the same expression under another name


def l2_loss(pred, gt, sqr_out=True, eps=1e-6):
    _check(pred, gt)
    return F.mse_loss(pred, gt)


def cosine_similarity_loss(pred, gt, eps=1e-6):
    _check(pred, gt)
    return 1 - F.cosine_similarity(gt.flatten(), pred.flatten(), dim=-1, eps=eps).mean()


def cg_alpha(r: torch.Tensor, d: torch.Tensor, q: torch.Tensor):
    """The first CG step length r.d / (d.q + 1e-6), loss.py:72-87."""
    assert r.shape == d.shape == q.shape
    return torch.dot(r.flatten(), d.flatten()) / (torch.dot(d.flatten(), q.flatten()) + 1e-6)


class NeuralCGLossBase(nn.Module):
    def __init__(self, batch_less: bool, block_size: int):
        super().__init__()
        self.spmv = GraphSpmv()
        self.spmv_transpose = GraphSpmv(True)
        self.block_size = block_size
        self.batch_less = batch_less

    def _ranges(self, batch) -> List[tuple]:
        if self.batch_less:
            ptrs = [0, int(batch.batch.numel())]
        else:
            ptrs = batch.ptr.detach().cpu().numpy().tolist()
        return list(zip(ptrs[:-1], ptrs[1:]))

    def _A(self, batch, d, mask=True):
        return self.spmv(X=d, edge_index=batch.edge_index, A=batch.matrix_values, mask=batch.mask if mask else None)

    def _mean_over_samples(self, batch, per_sample: Callable):
        ranges = self._ranges(batch)
        loss = 0.0
        for b, e in ranges:
            loss = loss + per_sample(b, e)
        return loss / len(ranges)

    def forward(self, batch, d: torch.Tensor, L_values: torch.Tensor):
        raise NotImplementedError()


class _PerSampleLoss(NeuralCGLossBase):
    def __init__(self, batch_less: bool, block_size: int, sqr_out: bool = True, eps: float = 1e-6):
        super().__init__(batch_less, block_size)
        self.sqr_out, self.eps = sqr_out, eps

    def fn(self, pred, gt):
        return rel_l2_loss(pred, gt, sqr_out=self.sqr_out, eps=self.eps)


class RelativeL2Loss_ANorm(_PerSampleLoss):
    """A d against r (loss.py:168-188) -- the default."""

    def forward(self, batch, d, L_values):
        Ad = self._A(batch, d)
        return self._mean_over_samples(batch, lambda b, e: self.fn(Ad[b:e], batch.residual[b:e]))


class RelativeL2Loss_PlainNorm(_PerSampleLoss):
    """d against the stored solution ``batch.gt`` (loss.py:142-165)."""

    def forward(self, batch, d, L_values):
        return self._mean_over_samples(batch, lambda b, e: self.fn(d[b:e], batch.gt[b:e]))


class NifLoss(_PerSampleLoss):
    """d against A r, unmasked product (loss.py:119-139)."""

    def forward(self, batch, d, L_values):
        Ar = self._A(batch, batch.residual, mask=False)
        return self._mean_over_samples(batch, lambda b, e: self.fn(d[b:e], Ar[b:e]))


class L2Loss_ANorm(_PerSampleLoss):
    def forward(self, batch, d, L_values):
        return F.mse_loss(self._A(batch, d), batch.residual)


class L1Loss(_PerSampleLoss):
    def forward(self, batch, d, L_values):
        return F.l1_loss(self._A(batch, d), batch.residual, reduction="mean")


class _CosineLoss(NeuralCGLossBase):
    def __init__(self, batch_less: bool, block_size: int, eps: float = 1e-6):
        super().__init__(batch_less, block_size)
        self.eps = eps

    def fn(self, pred, gt):
        return cosine_similarity_loss(pred, gt, eps=self.eps)


class CosineSimilarityLoss_PlainNorm(_CosineLoss):
    def forward(self, batch, d, L_values):
        return self._mean_over_samples(batch, lambda b, e: self.fn(d[b:e], batch.gt[b:e]))


class CosineSimilarityLoss_ANorm(_CosineLoss):
    def forward(self, batch, d, L_values):
        Ad = self._A(batch, d)
        return self._mean_over_samples(batch, lambda b, e: self.fn(Ad[b:e], batch.residual[b:e]))


class ConjGradLoss_PlainNorm(_PerSampleLoss):
    """One CG step x1 = alpha d against the stored solution (loss.py:244-277)."""

    def forward(self, batch, d, L_values):
        r, Ad = batch.residual, self._A(batch, d)
        return self._mean_over_samples(
            batch, lambda b, e: self.fn(cg_alpha(r[b:e], d[b:e], Ad[b:e]) * d[b:e], batch.gt[b:e]))


class ConjGradLoss_ANorm(_PerSampleLoss):
    """alpha A d against r (loss.py:280-308)."""

    def forward(self, batch, d, L_values):
        r, Ad = batch.residual, self._A(batch, d)
        return self._mean_over_samples(
            batch, lambda b, e: self.fn(cg_alpha(r[b:e], d[b:e], Ad[b:e]) * Ad[b:e], r[b:e]))


class ConjGradLoss_ANorm_NoRelative(ConjGradLoss_ANorm):
    def fn(self, pred, gt):
        return l2_loss(pred, gt)


class PropLoss(_PerSampleLoss):
    """loss.py:343-372: the per-sample projection residual is computed and discarded; the value is the MSE."""

    def forward(self, batch, d, L_values):
        return F.mse_loss(self._A(batch, d), batch.residual)


class RelPropLoss(PropLoss):
    pass


_LOSSES = {cls.__name__.lower(): cls for cls in (
    RelativeL2Loss_PlainNorm, RelativeL2Loss_ANorm, L2Loss_ANorm, PropLoss, L1Loss, RelPropLoss,
    CosineSimilarityLoss_PlainNorm, CosineSimilarityLoss_ANorm, ConjGradLoss_PlainNorm, ConjGradLoss_ANorm,
    ConjGradLoss_ANorm_NoRelative)}
_LOSSES["nifloss_norm"] = NifLoss


def create_loss_item(name: str, batch_less, block_size, **params) -> NeuralCGLossBase:
    try:
        cls = _LOSSES[name.lower()]
    except KeyError:
        raise ValueError(f"Unknown loss {name}") from None
    return cls(batch_less=batch_less, block_size=block_size, **params)
