"""Host-side mirror of the reference's ``neural_cg`` package for the learned-SPAI solve path.

Same module paths, class names, constructor signatures and ``state_dict`` keys as the reference
(SURVEY.md §8b); everything that computes calls libb200spai.so through ``ctypes``.  Importable as
``learningsparsepreconditioner4gpu_b200.neural_cg`` or, after ``install_as_neural_cg()``, as plain
``neural_cg`` so that the reference's own scripts (``infer.py``) pick it up unchanged.
"""
import sys as _sys


def install_as_neural_cg() -> None:
    """Alias this package (and its loaded sub-modules) as top-level ``neural_cg``."""
    me = _sys.modules[__name__]
    _sys.modules.setdefault("neural_cg", me)
    for name, mod in list(_sys.modules.items()):
        if name.startswith(__name__ + "."):
            _sys.modules.setdefault("neural_cg" + name[len(__name__):], mod)
