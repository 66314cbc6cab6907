"""Device time of one training step through the fused kernels (measurement, GPU box).

    python scripts/train_step_bench.py [grid=64] [steps=5]

One jittered Kuhn-tet Poisson system of grid^3 nodes as a one-sample batch, random-init weights, RelativeL2Loss_ANorm:
forward with tape -> AATPE -> loss -> backward (CUDA backward kernels) -> AdamW step; CUDA events around the phases."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from learningsparsepreconditioner4gpu_b200 import _lib, systems

    grid = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    s = systems.poisson_sample(grid, seed=5, gen_device="cuda", device="cuda")
    s.residual = torch.randn(s.num_nodes, 1, device="cuda") * s.mask
    s.batch = torch.zeros(s.num_nodes, dtype=torch.long, device="cuda")
    ws = systems.default_workspace(s).cuda().train()
    opt = torch.optim.AdamW(ws.parameters(), lr=1e-3, weight_decay=3e-3)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    acc, losses = [0.0, 0.0, 0.0], []
    l0 = 0
    for it in range(steps + 2):
        if it == 2:
            acc, l0 = [0.0, 0.0, 0.0], _lib.total_launches()
        opt.zero_grad(set_to_none=True)
        ev[0].record()
        loss = ws.training_step(s, 0)
        ev[1].record()
        loss.backward()
        ev[2].record()
        opt.step()
        ev[3].record()
        torch.cuda.synchronize()
        for k in range(3):
            acc[k] += ev[k].elapsed_time(ev[k + 1])
        losses.append(loss.item())
    print(json.dumps({"grid": grid, "N": s.num_nodes, "E": s.num_edges, "steps": steps,
                      "ms_forward_aatpe_loss": round(acc[0] / steps, 3), "ms_backward": round(acc[1] / steps, 3),
                      "ms_optimizer": round(acc[2] / steps, 3), "ms_step": round(sum(acc) / steps, 3),
                      "library_launches_per_step": (_lib.total_launches() - l0) / steps,
                      "loss_first_last": [losses[0], losses[-1]],
                      "peak_mem_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 2)}), flush=True)


if __name__ == "__main__":
    main()
