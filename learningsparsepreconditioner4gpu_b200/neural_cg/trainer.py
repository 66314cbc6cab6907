"""The fit loop the reference gets from ``L.Trainer(**cfg.trainer).fit(model, datamodule=data)`` (train.py:98-110),
reduced to what ``config/trainer.yaml`` uses: ``max_epochs``, ``accumulate_grad_batches``, ``gradient_clip_val`` (norm
clipping, Lightning's default algorithm), ``check_val_every_n_epoch``, an epoch-wise scheduler step, ``fast_dev_run``.

Lightning is absent from this image; this is host plumbing around ``workspace.training_step`` -- all arithmetic of a
step (GNN forward + tape, AATPE, loss product, both backward passes) runs in libb200spai.so, the optimiser update is
torch's.  Batches must live on the GPU (``Data.to("cuda")``).

Several GPUs (one process per GPU under torchrun, ``torch.distributed`` initialised by the caller): plain data parallelism --
``shard_batches`` deals the batches to the ranks, ``allreduce_mean_grads`` averages the gradient over the ranks (one
all-reduce of 70 KB per optimiser step, the path's only exchange), every rank applies the same update."""
from __future__ import annotations

from typing import Any, Dict, Iterable, List, Optional, Sequence

import torch


def _dist():
    """(torch.distributed, rank, world) when a process group is up, else (None, 0, 1)."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist, dist.get_rank(), dist.get_world_size()
    return None, 0, 1


def shard_batches(batches: Iterable, rank: int, world: int) -> List:
    """Rank r takes batches r, r + world, ... of the first (len // world) * world: every rank makes the same number of steps
    (a rank with one batch more would wait in the gradient all-reduce for ever)."""
    items = list(batches)
    usable = (len(items) // world) * world
    return items[rank:usable:world]


def allreduce_mean_grads(params: Sequence[torch.nn.Parameter]) -> None:
    """Data-parallel training: the gradient of the step is the mean over the ranks' batches.  ONE all-reduce of the flat
    gradient (the whole GNN is ~18 k parameters = 70 KB: launch latency, not bandwidth) -- the only exchange step of the
    training path; every rank then applies the identical update, so the replicas never drift.  No-op without a group."""
    dist, _, world = _dist()
    if dist is None:
        return
    grads = [p.grad for p in params if p.grad is not None]
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    flat /= world
    off = 0
    for g in grads:
        g.copy_(flat[off:off + g.numel()].view_as(g))
        off += g.numel()


class Trainer:
    def __init__(self, max_epochs: int = 1, accumulate_grad_batches: int = 1, gradient_clip_val: Optional[float] = None,
                 check_val_every_n_epoch: int = 1, fast_dev_run: bool = False, **_ignored: Any):
        self.max_epochs = 1 if fast_dev_run else int(max_epochs)
        self.accumulate = max(1, int(accumulate_grad_batches))
        self.clip = gradient_clip_val
        self.val_every = max(1, int(check_val_every_n_epoch))
        self.fast_dev_run = bool(fast_dev_run)
        self.history: List[Dict[str, float]] = []
        self.global_step = 0

    def fit(self, model, train_batches: Iterable, val_batches: Optional[Iterable] = None, device: str = "cuda"):
        cfg = model.configure_optimizers()
        opt, sched = (cfg["optimizer"], cfg.get("lr_scheduler")) if isinstance(cfg, dict) else (cfg, None)
        params = [p for p in model.parameters() if p.requires_grad]
        _, rank, world = _dist()
        for epoch in range(self.max_epochs):
            model.train()
            opt.zero_grad(set_to_none=True)
            total, count = 0.0, 0
            # under torch.distributed (one process per GPU, NCCL): the epoch's batches are dealt round-robin to the ranks and
            # the gradients are averaged over the ranks before every optimiser step
            epoch_batches = shard_batches(train_batches, rank, world) if world > 1 else train_batches
            for i, batch in enumerate(epoch_batches):
                batch = batch.to(device) if hasattr(batch, "to") else batch
                loss = model.training_step(batch, i)
                (loss / self.accumulate).backward()
                if (i + 1) % self.accumulate == 0:
                    allreduce_mean_grads(params)
                    if self.clip:
                        torch.nn.utils.clip_grad_norm_(params, float(self.clip))
                    opt.step()
                    opt.zero_grad(set_to_none=True)
                    self.global_step += 1
                total += float(loss.detach())
                count += 1
                if self.fast_dev_run:
                    break
            if sched is not None:
                sched.step()
            rec = {"epoch": epoch, "train_loss": total / max(count, 1)}
            if val_batches is not None and (epoch + 1) % self.val_every == 0:
                model.eval()
                vt, vc = 0.0, 0
                for i, batch in enumerate(val_batches):
                    batch = batch.to(device) if hasattr(batch, "to") else batch
                    with torch.no_grad():
                        out = model.validation_step(batch, i)
                    if torch.is_tensor(out) and out.dim() == 0:
                        vt += float(out)
                        vc += 1
                    if self.fast_dev_run:
                        break
                if vc:
                    rec["val_loss"] = vt / vc
            self.history.append(rec)
        return self.history
