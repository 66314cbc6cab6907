"""Mint tests/golden/train_losses.npz from the UNMODIFIED reference loss classes (run once in the build container).

    python tests/golden/make_golden_losses.py

Every name `create_loss_item` accepts (neural_cg/loss.py:420-459) is evaluated on one two-sample batch (the poisson pair of
train_poisson_pair_b1.npz, plus a seeded `gt` for the PlainNorm losses) at a seeded d = "M^-1 r": the loss value and its
gradient w.r.t. d through torch autograd.  Same shim as the other generators (GraphSpmv runs on the PyG stand-in)."""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_shim  # noqa: E402

ref_shim.install("/root/reference")

from neural_cg.loss import create_loss_item  # noqa: E402  (reference)

from conftest import load_train_case  # noqa: E402

NAMES = ["RelativeL2Loss_PlainNorm", "RelativeL2Loss_ANorm", "L2Loss_ANorm", "PropLoss", "L1Loss", "RelPropLoss",
         "CosineSimilarityLoss_PlainNorm", "CosineSimilarityLoss_ANorm", "ConjGradLoss_PlainNorm", "ConjGradLoss_ANorm",
         "ConjGradLoss_ANorm_NoRelative", "NifLoss_Norm"]


def main():
    b, _, _ = load_train_case("poisson_pair_b1")
    g = torch.Generator().manual_seed(21)
    n = b["x"].shape[0]
    d0 = torch.randn(n, 1, generator=g) * b["mask"]
    gt = torch.randn(n, 1, generator=g) * b["mask"]
    out = {"d": d0.numpy(), "gt": gt.numpy(), "names": np.array(NAMES)}
    for batch_less in (False, True):
        batch = ref_shim.Data(**b, gt=gt)
        for name in NAMES:
            d = d0.clone().requires_grad_(True)
            loss = create_loss_item(name, batch_less=batch_less, block_size=1)(batch, d, None)
            loss.backward()
            tag = f"{name}/{'batch_less' if batch_less else 'per_sample'}"
            out[tag + "/loss"] = np.float64(loss.item())
            out[tag + "/grad_d"] = d.grad.numpy().copy()
            print(f"{tag}: {loss.item():.7f}  |grad| {float(d.grad.norm()):.4e}")
    path = os.path.join(HERE, "train_losses.npz")
    np.savez_compressed(path, **out)
    print("->", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
