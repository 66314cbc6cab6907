"""Host-side mirror of ``neural_cg/utils/optim.py``: optimiser / scheduler factories by config name
(``config/trainer.yaml``: adamw + exp by default).  Unknown names raise ``ValueError`` like the reference."""
from __future__ import annotations

from torch import optim
from torch.optim import lr_scheduler

_OPTIMIZERS = {"adam": optim.Adam, "adamw": optim.AdamW, "sgd": optim.SGD}
_SCHEDULERS = {"onecycle": lr_scheduler.OneCycleLR, "exp": lr_scheduler.ExponentialLR,
               "cosine": lr_scheduler.CosineAnnealingLR, "cosine_warmup": lr_scheduler.CosineAnnealingWarmRestarts}


def create_optimizer(parameters, name, params):
    if name not in _OPTIMIZERS:
        raise ValueError(f"Unknown optimizer {name}")
    return _OPTIMIZERS[name](parameters, **dict(params or {}))


def create_scheduler(optimizer, name, params):
    if name == "none":
        return None
    if name not in _SCHEDULERS:
        raise ValueError(f"Unknown scheduler {name}")
    return _SCHEDULERS[name](optimizer, **dict(params or {}))
