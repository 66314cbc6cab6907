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


def solve_batch(make_workspace: Callable[[], object], samples: Sequence, streams: int = 4, **kw):
    """One-shot convenience around ``BatchSolver``."""
    bs = BatchSolver(make_workspace, streams)
    try:
        return bs.solve(samples, **kw)
    finally:
        bs.close()
