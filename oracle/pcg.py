"""ORACLE (test infrastructure, not product): CPU restatement of the reference's PCG solve.

The reference's only in-tree definition of the ``ext_spai`` solve is the scipy fallback
neural_cg/utils/validate.py:163-201 (get_pcg_iter_time_scipy): ``scipy.sparse.linalg.cg`` with
``M(x) = G (G^T x) + eps x`` where ``G^T`` is materialised once as CSR (:176,:182), ``b = A @ gt``,
``x0 = 0``, ``max_iter = n`` when 0, iteration count = number of callbacks.  The timed solver,
pymathprim (github.com/Adversarr/mathprim, unpinned, not in the tree), is opaque: parity with IT is
unpinned; parity with the scipy definition is pinned (tests/test_oracle_golden.py: iteration counts
and true-residual histories recorded from the reference function itself).

``pcg_restated`` writes the scipy 1.15/1.18 ``cg`` recurrence out (scipy/sparse/linalg/_isolve/
iterative.py) so that the residual history can be captured and so that the CUDA loop can be compared
step by step:

    atol = max(0, rtol * ||b||);  r = b - A x0 (= b);  for k in range(maxiter):
        if ||r|| < atol: return            # strict <, recurrence residual, tested at the loop top
        z = M r;  rho = r.z;  p = z (k == 0) else z + (rho/rho_prev) p
        q = A p;  alpha = rho / p.q;  x += alpha p;  r -= alpha q;  callback

Preconditioner variants: "ext_spai" (above), "ext_spai_scaled" (validate.py:267-302:
G((G^T x)/d) + eps x/d, d = diag(A)), "diagonal" (validate.py:235-264: x/d), "none".
"""
from __future__ import annotations

import time
from typing import Callable, List, Optional, Tuple

import numpy as np
from scipy.sparse import csr_matrix
from scipy.sparse.linalg import LinearOperator, cg


def make_preconditioner(kind: str, A: csr_matrix, G: Optional[csr_matrix], epsilon: float) -> Callable:
    if kind == "none":
        return lambda r: r.copy()
    if kind == "diagonal":
        d = A.diagonal()
        return lambda r: r / d
    GT = csr_matrix(G.T)
    if kind == "ext_spai":
        return lambda r: G @ (GT @ r) + epsilon * r
    if kind == "ext_spai_scaled":
        d = A.diagonal()
        return lambda r: G @ ((GT @ r) / d) + epsilon * r / d
    raise ValueError(f"Unknown preconditioner {kind}")


def pcg_restated(
    A: csr_matrix,
    b: np.ndarray,
    M: Callable,
    rtol: float = 1e-6,
    max_iter: int = 0,
) -> Tuple[np.ndarray, int, List[float]]:
    """Returns (x, iterations, history) with history[k] = ||r_k||_2 of the recurrence residual
    (history[0] = ||b||; len = iterations + 1)."""
    n = A.shape[0]
    max_iter = max_iter if max_iter > 0 else n
    b = np.asarray(b, dtype=np.float64)
    atol = max(0.0, rtol * float(np.linalg.norm(b)))
    x = np.zeros(n)
    r = b.copy()
    hist = [float(np.linalg.norm(r))]
    if hist[0] == 0.0:  # scipy returns immediately for b == 0
        return x, 0, hist
    rho_prev, p = None, None
    its = 0
    for k in range(max_iter):
        if hist[-1] < atol:
            break
        z = M(r)
        rho = float(r @ z)
        if k > 0:
            p *= rho / rho_prev
            p += z
        else:
            p = z.copy()
        q = A @ p
        alpha = rho / float(p @ q)
        x += alpha * p
        r -= alpha * q
        rho_prev = rho
        its += 1
        hist.append(float(np.linalg.norm(r)))
    return x, its, hist


def pcg_scipy(A, gt, G, epsilon, rtol=1e-6, max_iter=0, kind="ext_spai", with_time=False, history=False):
    """The reference's scipy solve (validate.py:163-201 and siblings): b = A @ gt, scipy cg, callback count.

    ``history=True`` additionally records the TRUE residual ||b - A x_k|| per callback."""
    n = A.shape[0]
    max_iter = max_iter if max_iter > 0 else n
    A = A.astype(np.float64)
    Mfun = make_preconditioner(kind, A, None if G is None else G.astype(np.float64), epsilon)
    M = LinearOperator((n, n), matvec=Mfun, dtype=np.float64)
    b = A @ gt
    count = 0
    true_res = []

    def cb(xk):
        nonlocal count
        count += 1
        if history:
            true_res.append(float(np.linalg.norm(b - A @ xk)))

    t0 = time.time()
    x, _ = cg(A, b, M=M, callback=cb, rtol=rtol, maxiter=max_iter)
    dt = time.time() - t0
    out = [count]
    if with_time:
        out.append(dt)
    if history:
        out.append(true_res)
        out.append(x)
    return out[0] if len(out) == 1 else tuple(out)


class ScipyPcgTimer:
    """The reference's scipy solve with the operator built ONCE (as validate.py:173-190 does before its clock starts),
    so that bench.py can time a bounded number of iterations per step without re-materialising G^T."""

    def __init__(self, A, G, epsilon, kind="ext_spai"):
        self.A = A.astype(np.float64)
        n = A.shape[0]
        self.M = LinearOperator((n, n), matvec=make_preconditioner(kind, self.A, None if G is None else G.astype(np.float64),
                                                                   epsilon), dtype=np.float64)

    def run(self, gt, rtol=1e-6, max_iter=0):
        """-> (callbacks, seconds): ``cg`` bracketed by the clock exactly like validate.py:195-197."""
        b = self.A @ gt
        count = 0

        def cb(xk):
            nonlocal count
            count += 1

        t0 = time.time()
        cg(self.A, b, M=self.M, callback=cb, rtol=rtol, maxiter=max_iter if max_iter > 0 else self.A.shape[0])
        return count, time.time() - t0
