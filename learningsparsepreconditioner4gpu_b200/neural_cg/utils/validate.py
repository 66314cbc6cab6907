"""Host-side mirror of ``neural_cg/utils/validate.py``: CSR assembly and the PCG harness, same names and
keyword sets, executed on the B200 through libb200spai.so.

* ``to_csr_cpu`` (:22-51) keeps its name and its return type (scipy CSR, fp64, int32 sorted indices) but
  the assembly -- block expansion, Dirichlet masking, exact-zero pruning -- runs in ``b200spai_edges_to_csr``;
  ``to_csr_device`` is the same without the download.
* ``get_pcg_iter_time`` (:89-121), ``get_pcg_scaled_iter_time`` (:124-160) and ``get_cg_iter_time``
  (:54-86) build ``linalg.PreconditionedConjugateGradient`` exactly where the reference builds pymathprim's.
  ``device`` defaults to "cuda"; "cpu" raises (there is no CPU path).  ``get_cg_iter_time`` knows the
  methods "none" and "diagonal" (ainv / ic / fsai are comparison baselines outside this path).
* ``get_pcg_iter_time_scipy`` (:163-201) keeps the name of the reference's pymathprim-less fallback and
  its return value (iteration count [, seconds]); the solve is the same device solve.
"""
from __future__ import annotations

import time
from typing import Optional, Tuple, Union

import numpy as np
import torch
from scipy.sparse import csr_matrix

from ... import _lib
from ...device import DeviceCSR
from ...linalg import PreconditionedConjugateGradient as pcg

MatrixLike = Union[csr_matrix, DeviceCSR]


def to_numpy(x) -> np.ndarray:
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    if isinstance(x, np.ndarray):
        return x
    raise ValueError(f"Unknown type {type(x)}")


def _cuda(x, dtype=None) -> Optional[torch.Tensor]:
    if x is None:
        return None
    t = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x))
    return t.to(device="cuda", dtype=dtype) if dtype is not None else t.to(device="cuda")


def to_csr_device(edge_index, edge_attr, n: int, mask, dtype=np.float64, plan: bool = False) -> DeviceCSR:
    """Edge list (+ mask) -> fp64 CSR resident in HBM; semantics of to_csr_cpu.

    Restriction against the reference (validate.py:22-51 widens WHATEVER dtype arrives to float64): the assembly entry
    point takes fp32 edge values -- what every caller on the path passes (``matrix_values`` and the GNN output are fp32
    tensors, data.py:312, workspace.py:94).  float64 values that are not exactly fp32-representable would silently
    lose bits, so they are refused; build such a matrix with ``DeviceCSR.from_scipy`` (fp64 values) instead."""
    _lib.require_cuda()
    if np.dtype(dtype) != np.float64:
        raise ValueError("only float64 matrices are supported (the reference assembles in float64)")
    assert edge_index.ndim == 2 and edge_index.shape[0] == 2
    assert edge_attr.ndim in (1, 3)
    vals = _cuda(edge_attr, torch.float32)
    wide = edge_attr.dtype in (torch.float64, np.float64) if hasattr(edge_attr, "dtype") else False
    if wide and not bool((vals.to(torch.float64) == _cuda(edge_attr, torch.float64)).all()):
        raise ValueError("to_csr: float64 edge values that are not fp32-representable would lose precision in the fp32 "
                         "assembly entry point; pass fp32 values or build the matrix with DeviceCSR.from_scipy")
    bsize = 1 if vals.dim() == 1 else vals.shape[-1]
    assert n % bsize == 0
    return DeviceCSR.from_edges(_cuda(edge_index, torch.int64), vals, n // bsize, _cuda(mask, torch.float32), plan=plan)


def to_csr_cpu(edge_index, edge_attr, n: int, mask, dtype=np.float64) -> csr_matrix:
    return to_csr_device(edge_index, edge_attr, n, mask, dtype).to_scipy()


def _check_device(device) -> None:
    if str(device) != "cuda":
        raise ValueError(f"device={device!r}: this implementation only solves on 'cuda' (B200); there is no CPU fallback")


def _rhs(A: MatrixLike, gt) -> np.ndarray:
    """b = A @ gt (validate.py:102)."""
    gt = np.ascontiguousarray(to_numpy(gt), dtype=np.float64).reshape(-1)
    if isinstance(A, DeviceCSR):
        return A.matvec(torch.from_numpy(gt).cuda()).cpu().numpy()
    return np.asarray(A @ gt, dtype=np.float64).copy()


def _run(A: MatrixLike, gt, preconditioner: str, rtol, max_iter, repeat, device, ext_spai=None,
         raise_on_max_iter: bool = False) -> Tuple[float, float, float]:
    _check_device(device)
    assert repeat > 0
    rows = A.shape[0]
    max_iter = max_iter if max_iter > 0 else rows
    b = _rhs(A, gt)
    if not isinstance(A, DeviceCSR):
        A = DeviceCSR.from_scipy(csr_matrix(A).astype(np.float64))
    if ext_spai is not None and not isinstance(ext_spai[0], DeviceCSR):
        ext_spai = (DeviceCSR.from_scipy(csr_matrix(ext_spai[0]).astype(np.float64)), ext_spai[1])
    iter_cnt, time_prec, time_elp = 0, 0.0, 0.0
    x = np.zeros_like(b)
    for _ in range(repeat):
        x_copy, b_copy = x.copy(), b.copy()
        solver = pcg(A, device=device, preconditioner=preconditioner, dtype=np.float64)
        kw = {} if ext_spai is None else {"ext_spai": ext_spai}
        this_iter, this_prec, this_solve = solver(b_copy, x_copy, rtol, max_iter, **kw)
        iter_cnt += this_iter
        time_prec += this_prec
        time_elp += this_solve
        if raise_on_max_iter and this_iter >= max_iter:
            raise RuntimeError("CG did not converge")
    return iter_cnt / repeat, time_prec / repeat, time_elp / repeat


def get_cg_iter_time(A: MatrixLike, gt, rtol=1e-6, max_iter=0, dtype=np.float64, repeat=1, device="cuda",
                     method="diagonal") -> Tuple[float, float, float]:
    if method not in ("none", "diagonal"):
        raise ValueError(f"Unknown method {method}: the B200 path provides 'none' and 'diagonal' "
                         "(ainv / ic / fsai are comparison baselines of the reference, not part of this path)")
    return _run(A, gt, method, rtol, max_iter, repeat, device, raise_on_max_iter=True)


def get_pcg_iter_time(A: MatrixLike, gt, spai: MatrixLike, epsilon: float, rtol=1e-6, max_iter=0, repeat=1,
                      dtype=np.float64, device="cuda") -> Tuple[float, float, float]:
    return _run(A, gt, "ext_spai", rtol, max_iter, repeat, device, ext_spai=(spai, epsilon))


def get_pcg_scaled_iter_time(A: MatrixLike, gt, spai: MatrixLike, epsilon: float, rtol=1e-6, max_iter=0, repeat=1,
                             dtype=np.float64, device="cuda") -> Tuple[float, float, float]:
    return _run(A, gt, "ext_spai_scaled", rtol, max_iter, repeat, device, ext_spai=(spai, epsilon))


def _count(A, gt, kind, spai, epsilon, max_iter, rtol, with_time):
    t0 = time.time()
    iters, _, _ = _run(A, gt, kind, rtol, max_iter, 1, "cuda", ext_spai=None if spai is None else (spai, epsilon))
    dt = time.time() - t0
    return (int(iters), dt) if with_time else int(iters)


def get_pcg_iter_time_scipy(A: MatrixLike, gt, spai: MatrixLike, epsilon: float, max_iter=0, rtol=1e-6,
                            dtype=np.float64, with_time: bool = False):
    return _count(A, gt, "ext_spai", spai, epsilon, max_iter, rtol, with_time)


def get_pcg_scaled_iter_time_scipy(A: MatrixLike, gt, spai: MatrixLike, epsilon: float, max_iter=0, rtol=1e-6,
                                   dtype=np.float64, with_time: bool = False):
    return _count(A, gt, "ext_spai_scaled", spai, epsilon, max_iter, rtol, with_time)


def get_cg_iter_time_scipy(A: MatrixLike, gt, max_iter=0, rtol=1e-6, dtype=np.float64, with_time: bool = False):
    return _count(A, gt, "none", None, 0.0, max_iter, rtol, with_time)


def get_pcg_diagonal_iter_time_scipy(A: MatrixLike, gt, max_iter=0, rtol=1e-6, dtype=np.float64, with_time: bool = False):
    return _count(A, gt, "diagonal", None, 0.0, max_iter, rtol, with_time)


def get_pcg_stat_scipy(A: MatrixLike, r, prefix: str) -> dict:
    """Baseline iteration counts logged next to the neural one (validate.py:304-313, minus ic0)."""
    return {f"{prefix}/cuda_none_iter": get_cg_iter_time_scipy(A, r),
            f"{prefix}/cuda_diag_iter": get_pcg_diagonal_iter_time_scipy(A, r)}
