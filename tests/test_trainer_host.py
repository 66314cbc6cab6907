"""Host logic of the training loop stand-in (neural_cg/trainer.py; the reference: L.Trainer(**cfg.trainer).fit, train.py:98-110)
on the CPU: gradient clipping / accumulation / scheduler, and -- over a real 2-process gloo group -- the data-parallel path
(batches dealt to the ranks, ONE all-reduce of the flat gradient per optimiser step, identical replicas).

A toy model with the workspace's training interface (training_step / validation_step / configure_optimizers) stands in for the
CUDA workspace: the Trainer is plumbing and never looks inside a step.  tests/test_training_gpu.py runs it over the real one."""
import os

import numpy as np
import pytest
import torch
from torch import nn

from learningsparsepreconditioner4gpu_b200.neural_cg.trainer import Trainer, allreduce_mean_grads, shard_batches


class Toy(nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(0)
        self.net = nn.Sequential(nn.Linear(3, 5), nn.Tanh(), nn.Linear(5, 1))

    def training_step(self, batch, i):
        x, y = batch
        return ((self.net(x) - y) ** 2).mean()

    def validation_step(self, batch, i):
        return self.training_step(batch, i).detach()

    def configure_optimizers(self):
        opt = torch.optim.AdamW(self.parameters(), lr=1e-2, weight_decay=3e-3)
        return {"optimizer": opt, "lr_scheduler": torch.optim.lr_scheduler.ExponentialLR(opt, gamma=0.9)}


def _batches(n, seed=1):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn(8, 3, generator=g), torch.randn(8, 1, generator=g)) for _ in range(n)]


def _manual(batches, epochs, group, clip):
    """What fit() must compute: per optimiser step the MEAN gradient of a group of batches, clipped, AdamW, epoch-wise decay."""
    m = Toy()
    cfg = m.configure_optimizers()
    opt, sched = cfg["optimizer"], cfg["lr_scheduler"]
    params = list(m.parameters())
    for _ in range(epochs):
        usable = (len(batches) // group) * group
        for s in range(0, usable, group):
            opt.zero_grad(set_to_none=True)
            for b in batches[s:s + group]:
                (m.training_step(b, 0) / group).backward()
            torch.nn.utils.clip_grad_norm_(params, clip)
            opt.step()
        sched.step()
    return torch.cat([p.detach().reshape(-1) for p in params])


def test_fit_single_process_matches_manual_loop():
    batches = _batches(6)
    m = Toy()
    hist = Trainer(max_epochs=3, gradient_clip_val=0.5, accumulate_grad_batches=2, check_val_every_n_epoch=1).fit(
        m, batches, val_batches=batches[:2], device="cpu")
    got = torch.cat([p.detach().reshape(-1) for p in m.parameters()])
    assert torch.allclose(got, _manual(batches, 3, 2, 0.5), rtol=0, atol=1e-7)
    assert len(hist) == 3 and all("val_loss" in h for h in hist) and hist[-1]["train_loss"] < hist[0]["train_loss"]
    assert len(Trainer(max_epochs=7, fast_dev_run=True).fit(Toy(), batches, device="cpu")) == 1


def test_shard_batches_gives_every_rank_the_same_number():
    items = list(range(7))
    parts = [shard_batches(items, r, 3) for r in range(3)]
    assert parts == [[0, 3], [1, 4], [2, 5]]            # the 7th batch is dropped: nobody waits in an all-reduce
    assert shard_batches(items, 0, 1) == items
    allreduce_mean_grads(list(Toy().parameters()))       # no process group: a no-op


def _dp_worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        m = Toy()
        Trainer(max_epochs=3, gradient_clip_val=0.5).fit(m, _batches(7), device="cpu")   # 7 batches: one is dropped
        q.put((rank, torch.cat([p.detach().reshape(-1) for p in m.parameters()]).numpy()))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_data_parallel_fit_matches_the_mean_gradient_loop():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + ((os.getpid() + 77) % 2000)
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(res[0], res[1])                       # identical replicas, bit for bit
    # = one process whose step gradient is the mean over the two ranks' batches (batches 2k and 2k + 1)
    want = _manual(_batches(7), 3, 2, 0.5).numpy()
    assert np.allclose(res[0], want, rtol=0, atol=1e-6)
