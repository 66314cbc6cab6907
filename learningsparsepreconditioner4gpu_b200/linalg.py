"""Operator-level plugin: the solver object the reference obtains from ``pymathprim.linalg``.

Reference call sites (neural_cg/utils/validate.py:73-80,110-117,145-156):

    solver = PreconditionedConjugateGradient(A, device="cuda", preconditioner="ext_spai", dtype=np.float64)
    iters, t_prec, t_solve = solver(b, x, rtol, max_iter, ext_spai=(G_csr, epsilon))     # x updated in place

pymathprim is not in the reference tree and is unpinned; the semantics implemented are those of the
reference's in-tree scipy definition (validate.py:163-201): x0 = 0, stop when ||r|| < rtol*||b||,
``iters`` = completed updates.  Times are seconds (Python floats): ``t_prec`` = building the
preconditioner on the device (upload of G + G^T), ``t_solve`` = the device solve.

Only device="cuda" exists here -- the library has no CPU path and says so.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from .device import PRECOND, DeviceCSR, DevicePCG

SUPPORTED = tuple(PRECOND.keys())


class PreconditionedConjugateGradient:
    def __init__(self, matrix, device: str = "cuda", preconditioner: str = "ext_spai", dtype=np.float64, mode: int = 0):
        if preconditioner not in PRECOND:
            raise ValueError(f"Unknown preconditioner {preconditioner}; supported: {SUPPORTED}")
        if str(device) != "cuda":
            raise ValueError("learningsparsepreconditioner4gpu_b200 only solves on device='cuda' (no CPU fallback)")
        if np.dtype(dtype) != np.float64:
            raise ValueError("only float64 is supported (the reference solves in float64, validate.py:116)")
        self.preconditioner = preconditioner
        self.A = matrix if isinstance(matrix, DeviceCSR) else DeviceCSR.from_scipy(matrix)
        self.solver = DevicePCG(self.A, mode=mode)
        self.last_result = None
        self.history: Optional[np.ndarray] = None

    def __call__(self, b: np.ndarray, x: np.ndarray, rtol: float = 1e-6, max_iter: int = 0, ext_spai=None,
                 record_history: bool = False) -> Tuple[int, float, float]:
        t_prec = 0.0
        if self.preconditioner in ("ext_spai", "ext_spai_scaled"):
            if ext_spai is None:
                raise ValueError("ext_spai=(G_csr, epsilon) is required for the ext_spai preconditioners")
            G, eps = ext_spai
            Gd = G if isinstance(G, DeviceCSR) else DeviceCSR.from_scipy(G)
            import time

            import torch

            t0 = time.perf_counter()
            self.solver.set_spai(Gd, eps)
            torch.cuda.current_stream().synchronize()
            t_prec = time.perf_counter() - t0
        b = np.ascontiguousarray(b, dtype=np.float64)
        if not (isinstance(x, np.ndarray) and x.dtype == np.float64 and x.flags["C_CONTIGUOUS"]):
            raise ValueError("x must be a contiguous float64 numpy array (updated in place)")
        hist = None
        if record_history:
            hist = np.zeros((max_iter if max_iter > 0 else b.shape[0]) + 1, dtype=np.float64)
        res = self.solver.solve_host(b, x, self.preconditioner, rtol, max_iter, hist)
        self.last_result = res
        self.history = None if hist is None else hist[: res.iterations + 1]
        return int(res.iterations), float(t_prec), float(res.solve_ms) * 1e-3
