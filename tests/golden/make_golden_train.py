"""Mint the TRAINING fixtures tests/golden/train_*.npz from the UNMODIFIED reference (run once in the build container;
/root/reference does not exist on the GPU box).

    python tests/golden/make_golden_train.py

For every inference fixture (tests/golden/<case>.npz: inputs minted by make_golden.py) the reference's own
``SimpleTrainingWorkspace._shared_step`` (neural_cg/workspace.py:96-112: GNN forward -> AATPE -> loss) and
``ScaledTrainingWorkspace._shared_step`` (neural_cg/scaled_workspace.py:96-112, diag = inv_diag) are run on a one-sample
batch with ``RelativeL2Loss_ANorm`` (neural_cg/loss.py:168-188, the config default), then ``loss.backward()`` through
torch autograd.  One more fixture batches two systems (block-diagonal, ``ptr`` = [0, N1, N1 + N2]) to pin the
per-sample mean.  Recorded: the residual (seeded), the state_dict actually used (the fixture's weights with seeded noise
on the LayerNorm affines, which random init leaves at 1 / 0), the loss and the gradient of EVERY parameter.

Same shim as make_golden.py: only PyG's MessagePassing glue and the Lightning base class are stand-ins (ref_shim.py);
the GNN, AATPE, GraphSpmv, the loss and the workspace code are the reference's files executed as they are."""
from __future__ import annotations

import os
import sys
from types import SimpleNamespace

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import ref_shim  # noqa: E402

ref_shim.install("/root/reference")

from neural_cg.scaled_workspace import ScaledTrainingWorkspace  # noqa: E402  (reference)
from neural_cg.workspace import SimpleTrainingWorkspace  # noqa: E402  (reference)

from make_golden import EPS, GNN_CFG  # noqa: E402

CASES = ["poisson_b1", "heat_b1", "elast_b3", "synthetic_mild_b1"]


def build(cls, f_node, f_edge, block_size):
    return cls(batch_size=1, batch_less=False, block_size=block_size, test_max_iter=1000,
               optimizer=SimpleNamespace(name="adamw", params={}), scheduler=SimpleNamespace(name="exp", params={}),
               loss=SimpleNamespace(name="RelativeL2Loss_ANorm", params=None), inspect_norms=False,
               check_converge=False, check_methods=[], check_devices=[],
               node_features=f_node, edge_features=f_edge, gnn=GNN_CFG, epsilon=EPS).train()


def sample_of(g, seed):
    n, b = g["x"].shape[0], int(g["block_size"])
    res = torch.randn(n, b, generator=torch.Generator().manual_seed(seed)) * torch.from_numpy(g["mask"])
    return dict(x=torch.from_numpy(g["x"]), edge_index=torch.from_numpy(g["raw_edge_index"]),
                edge_attr=torch.from_numpy(g["edge_attr"]), mask=torch.from_numpy(g["mask"]),
                matrix_values=torch.from_numpy(g["matrix_values"]), inv_diag=torch.from_numpy(g["inv_diag"]), residual=res)


def collate(samples):
    offs = np.cumsum([0] + [s["x"].shape[0] for s in samples])
    out = {k: torch.cat([s[k] for s in samples], dim=0) for k in samples[0] if k != "edge_index"}
    out["edge_index"] = torch.cat([s["edge_index"] + int(o) for s, o in zip(samples, offs[:-1])], dim=1)
    out["ptr"] = torch.as_tensor(offs, dtype=torch.long)
    out["batch"] = torch.repeat_interleave(torch.arange(len(samples)), torch.as_tensor(np.diff(offs)))
    return out


def run(name, goldens, seed, cases=None):
    g0 = goldens[0]
    b = int(g0["block_size"])
    fields = collate([sample_of(g, seed + i) for i, g in enumerate(goldens)])
    batch = ref_shim.Data(**fields)
    sd = {str(k): torch.from_numpy(g0["sd/" + str(k)].copy()) for k in g0["sd_keys"]}
    gen = torch.Generator().manual_seed(seed + 100)
    for k, v in sd.items():
        if "pre_norm" in k:
            v.add_(0.3 * torch.randn(v.shape, generator=gen))
    out = dict(block_size=np.int64(b), epsilon=np.float64(EPS), sd_keys=np.array(list(sd.keys())),
               cases=np.array([str(c) for c in (cases or [name])]), residual=fields["residual"].numpy(), ptr=fields["ptr"].numpy())
    for k, v in sd.items():
        out["sd/" + k] = v.numpy()
    for tag, cls in (("simple", SimpleTrainingWorkspace), ("scaled", ScaledTrainingWorkspace)):
        ws = build(cls, fields["x"].shape[1], fields["edge_attr"].shape[1], b)
        ws.load_state_dict(sd)
        boo, loss = ws._shared_step(batch, "Train", 0)
        loss.backward()
        out[f"{tag}/loss"] = np.float64(loss.item())
        out[f"{tag}/G_l2"] = np.float64(boo.detach().double().norm().item())
        for k, p in ws.named_parameters():
            if p.requires_grad:
                assert p.grad is not None, k
                out[f"{tag}/grad/{k}"] = p.grad.numpy().copy()
        gmax = max(float(p.grad.abs().max()) for p in ws.parameters() if p.grad is not None)
        print(f"train_{name} [{tag}]: loss {loss.item():.6f}  max|grad| {gmax:.3e}")
    path = os.path.join(HERE, f"train_{name}.npz")
    np.savez_compressed(path, **out)
    print(f"  -> {path} ({os.path.getsize(path) / 1024:.0f} KiB)")


def main():
    def load(c):
        return np.load(os.path.join(HERE, c + ".npz"), allow_pickle=False)

    for i, c in enumerate(CASES):
        run(c, [load(c)], seed=11 + i)
    run("poisson_pair_b1", [load("poisson_b1"), load("poisson_zeros_b1")], seed=31, cases=["poisson_b1", "poisson_zeros_b1"])


if __name__ == "__main__":
    main()
