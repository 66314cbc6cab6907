"""Worker of the multi-GPU data-parallel training test: run under torchrun, one rank per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29651 \
        tests/dist_train_worker.py

Every rank builds the same 2 x world one-sample batches and the same random-init workspace; Trainer.fit deals the batches to
the ranks and averages the gradient over NCCL before every optimiser step.  Rank 0 also trains ONE process over all batches
with accumulate_grad_batches = world (the same mean gradient per step) and compares; every rank's weights must be identical."""
import json
import os
import sys
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from learningsparsepreconditioner4gpu_b200 import systems  # noqa: E402
from learningsparsepreconditioner4gpu_b200.neural_cg.data import collate  # noqa: E402
from learningsparsepreconditioner4gpu_b200.neural_cg.trainer import Trainer  # noqa: E402


def flat(ws):
    return torch.cat([p.detach().reshape(-1) for p in ws.parameters() if p.requires_grad])


def make_ws(sample):
    ws = systems.default_workspace(sample).cuda().train()
    ws.optimizer_cfg = SimpleNamespace(name="adamw", params={"lr": 1e-3, "weight_decay": 3e-3})
    ws.scheduler_cfg = SimpleNamespace(name="exp", params={"gamma": 0.99})
    return ws


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    g = torch.Generator().manual_seed(4)
    batches = []
    for k in range(2 * world):
        s = systems.poisson_sample(10 + (k % 3), seed=60 + k, device="cpu")
        s.residual = torch.randn(s.num_nodes, 1, generator=g) * s.mask
        batches.append(collate([s]).to("cuda"))
    ws = make_ws(batches[0])
    hist = Trainer(max_epochs=4, gradient_clip_val=10).fit(ws, batches)
    mine = flat(ws)
    every = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    identical = all(torch.equal(every[0], e) for e in every)
    out = {"rank": rank, "world": world, "identical_replicas": identical, "steps": len(hist), "loss": [h["train_loss"] for h in hist]}
    if rank == 0:
        # the same mean gradient per step from one process: accumulate over `world` consecutive batches
        ref = make_ws(batches[0])
        import learningsparsepreconditioner4gpu_b200.neural_cg.trainer as tr

        saved = tr._dist
        tr._dist = lambda: (None, 0, 1)          # a single-process fit inside the live group
        try:
            Trainer(max_epochs=4, gradient_clip_val=10, accumulate_grad_batches=world).fit(ref, batches)
        finally:
            tr._dist = saved
        dev = float((flat(ref) - mine).abs().max() / flat(ref).abs().max())
        out["max_rel_dev_vs_single_process"] = dev
        assert dev < 1e-5, dev
    assert identical
    dist.barrier()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
