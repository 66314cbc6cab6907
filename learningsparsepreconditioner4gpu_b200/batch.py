"""Batches of independent systems on one GPU (BASELINE config C2; the reference solves batch members one by one,
neural_cg/workspace.py:137-143).

A small system cannot fill a B200: its solve is a chain of latencies (small kernels, grid barriers, host round trips).
``solve_batch`` therefore keeps several systems in flight: K host threads, each with its own workspace (own GNN engine
and topology cache) and its own non-blocking CUDA stream.  libb200spai handles are independent of each other and ctypes
releases the GIL during the calls, so the chains of different systems overlap.  Every system goes through exactly the
same kernels as in a sequential run; results are bit-identical to it (tests/test_pcg_gpu.py).
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Optional, Sequence, Tuple

import torch

from . import _lib


class BatchSolver:
    """``make_workspace()`` must return a NEW workspace each time (same weights: e.g. build + ``load_state_dict``).
    Collective resources (threads, streams, workspaces) are created once and reused by every ``solve``."""

    def __init__(self, make_workspace: Callable[[], object], streams: int = 4):
        _lib.require_cuda()
        self.k = max(int(streams), 1)
        self.workspaces = [make_workspace() for _ in range(self.k)]
        self.streams = [torch.cuda.Stream() for _ in range(self.k)]
        self.pool = ThreadPoolExecutor(max_workers=self.k) if self.k > 1 else None
        self.device = torch.cuda.current_device()

    def solve(self, samples: Sequence, rtol: float = 1e-6, max_iter: int = 0, preconditioner: str = "ext_spai",
              rhs: Optional[Sequence[Optional[torch.Tensor]]] = None) -> List[Tuple[torch.Tensor, object]]:
        """[(x, PcgResult)] in the order of ``samples`` (device-resident ``Data`` objects; ``rhs[j]`` or b = A @ mask)."""
        out: List[Optional[Tuple[torch.Tensor, object]]] = [None] * len(samples)
        torch.cuda.current_stream().synchronize()            # the inputs were produced on the caller's stream

        def worker(k: int) -> None:
            torch.cuda.set_device(self.device)
            with torch.cuda.stream(self.streams[k]):
                for j in range(k, len(samples), self.k):
                    x, res, _, A, G = self.workspaces[k].solve(samples[j], rhs=None if rhs is None else rhs[j], rtol=rtol,
                                                               max_iter=max_iter, preconditioner=preconditioner)
                    out[j] = (x, res)
                    del A, G
                self.streams[k].synchronize()

        if self.pool is None:
            worker(0)
        else:
            list(self.pool.map(worker, range(self.k)))
        return out  # type: ignore[return-value]

    def close(self):
        if self.pool is not None:
            self.pool.shutdown()
            self.pool = None


class CollatedBatchSolver:
    """The whole batch as ONE block-diagonal system, the way the reference's DataLoader collates it (``data.collate``:
    node / edge tensors concatenated, ``edge_index`` with cumulative node offsets, ``ptr``): one graph preprocessing, one
    GNN forward (10 launches for all systems), one assembly of A and G, one G^T -- and then every block solved as an
    independent system by its own CTA group inside one launch per wave (``DevicePCG.solve_batch`` ->
    ``b200spai_pcg_solve_batch``), each with its own scalars, stopping rule and iteration count, as
    ``log_converge_batch`` (workspace.py:137-143) does one by one.  No host thread per system, ~20 kernel launches per
    batch instead of ~17 per system."""

    def __init__(self, workspace):
        _lib.require_cuda()
        self.ws = workspace

    @torch.no_grad()
    def solve(self, samples: Sequence, rtol: float = 1e-6, max_iter: int = 0, preconditioner: str = "ext_spai",
              history: bool = False):
        """[(x, PcgResult)] in the order of ``samples`` (device-resident ``Data`` objects, b = A @ mask); with
        ``history`` the tuples carry a third entry, the residual history of that system."""
        from .device import DevicePCG
        from .neural_cg.data import collate
        from .neural_cg.utils.validate import to_csr_device

        batch = collate(samples)
        bsz = self.ws.block_size
        n = batch.x.shape[0] * bsz
        A = to_csr_device(batch.edge_index, batch.matrix_values, n, batch.mask)
        G = self.ws.inference_step_device(batch)
        rhs = A.matvec(batch.mask.reshape(-1).to(torch.float64))
        solver = DevicePCG(A)
        solver.set_spai(G, self.ws.epsilon)
        ptr = (batch.ptr.detach().cpu().numpy().astype("int64") * bsz)
        x, results, hists = solver.solve_batch(ptr, rhs, preconditioner, rtol=rtol, max_iter=max_iter, history=history)
        out = []
        for i in range(len(samples)):
            xi = x[int(ptr[i]): int(ptr[i + 1])]
            out.append((xi, results[i], hists[i]) if history else (xi, results[i]))
        self.last = (A, G, solver)     # keeps the handles alive until the caller has consumed x
        return out


def solve_batch(make_workspace: Callable[[], object], samples: Sequence, streams: int = 4, **kw):
    """One-shot convenience around ``BatchSolver``."""
    bs = BatchSolver(make_workspace, streams)
    try:
        return bs.solve(samples, **kw)
    finally:
        bs.close()
