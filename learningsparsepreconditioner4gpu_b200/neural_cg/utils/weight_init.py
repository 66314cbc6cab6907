"""neural_cg/utils/weight_init.py:1-4 -- re-draw every sub-module that knows how to."""
from torch import nn


def weight_init(m: nn.Module) -> None:
    reset = getattr(m, "reset_parameters", None)
    if callable(reset):
        reset()
