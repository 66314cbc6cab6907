"""Row-partitioned solve of ONE large system over the GPUs of a box (BASELINE config C5).

The reference is single-process; this is the new multi-GPU mode the north star asks for.  One process per GPU
(``torchrun``); ``torch.distributed`` is used only as plumbing -- to agree on the layout and to exchange the CUDA-IPC
handles of the peer-visible vector segments.  Inside the solve there is no NCCL call and no host involvement: halos are
pushed over NVLink by peer stores and the CG scalars are all-reduced through peer memory in rank order
(``csrc/dist.cu``), so every rank computes bit-identical alpha / beta / iteration counts.

Host-side pieces (pure index arithmetic, testable on CPU):

* ``partition_rows``      contiguous row blocks balanced on a per-row weight (nnz_A + nnz_G + nnz_GT);
* ``extended_range``      the column window a rank's rows touch (owned rows + halo), as a contiguous range;
* ``rank_layout``         (row_start, row_end, shift, hi): the extended vector of a rank holds global index c at c - shift;
* ``halo_plan``           who pushes which slice to whom (mirror of ``b200spai_dist_connect``).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from .device import PRECOND, DeviceCSR, _ptr, _stream

Layout = Tuple[int, int, int, int]  # row_start, row_end, shift, hi


# ---------------------------------------------------------------------------------------------------------
# host-side index arithmetic
# ---------------------------------------------------------------------------------------------------------
def partition_rows(row_weight: torch.Tensor, world: int) -> List[int]:
    """Contiguous row blocks with (nearly) equal total weight.  Deterministic integer arithmetic:
    row_start[k] = number of rows whose inclusive weight prefix is <= floor(k * total / world)."""
    n = int(row_weight.numel())
    c = torch.cumsum(row_weight.to(torch.int64), 0)
    total = int(c[-1].item()) if n else 0
    targets = torch.tensor([(k * total) // world for k in range(1, world)], dtype=torch.int64, device=c.device)
    cuts = torch.searchsorted(c, targets, right=True).tolist() if world > 1 else []
    starts = [0] + [int(v) for v in cuts] + [n]
    for k in range(1, len(starts)):  # monotone even with zero-weight rows
        starts[k] = max(starts[k], starts[k - 1])
    return starts


def extended_range(edge_index: torch.Tensor, rs: int, re: int) -> Tuple[int, int]:
    """Smallest contiguous index range [lo, hi) containing the owned rows [rs, re) and every column that rows
    [rs, re) of the matrix -- or of its transpose -- touch.  ``edge_index`` is the reference's [2, E] (row, col) list."""
    row, col = edge_index[0], edge_index[1]
    lo, hi = rs, re
    m = (row >= rs) & (row < re)
    if bool(m.any()):
        c = col[m]
        lo, hi = min(lo, int(c.min())), max(hi, int(c.max()) + 1)
    m = (col >= rs) & (col < re)  # rows of the transpose
    if bool(m.any()):
        r = row[m]
        lo, hi = min(lo, int(r.min())), max(hi, int(r.max()) + 1)
    return lo, hi


def rank_layout(rs: int, re: int, lo: int, hi: int) -> Layout:
    """shift = lo, moved down by one when needed so that the owned slice starts at an even extended index
    (the vector kernels use 16-byte accesses)."""
    shift = lo - ((rs - lo) & 1)
    return (rs, re, shift, hi)


def halo_plan(layouts: Sequence[Layout], rank: int):
    """(pushes, recvs): pushes = [(peer, src_off, dst_off, length)] -- owned entries of ``rank`` that ``peer``'s
    extended range contains; recvs = [peer] whose owned rows intersect this rank's extended range."""
    rs, re, shift, hi = layouts[rank]
    pushes, recvs = [], []
    for q, (rs_q, re_q, lo_q, hi_q) in enumerate(layouts):
        if q == rank:
            continue
        a, b = max(rs, lo_q), min(re, hi_q)
        if a < b:
            pushes.append((q, a - shift, a - lo_q, b - a))
        a2, b2 = max(rs_q, shift), min(re_q, hi)
        if a2 < b2:
            recvs.append(q)
    return pushes, recvs


# ---------------------------------------------------------------------------------------------------------
# the distributed solver
# ---------------------------------------------------------------------------------------------------------
def _all_gather_object(obj, group):
    import torch.distributed as dist

    out = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, obj, group=group)
    return out


class DistPCG:
    """Row-partitioned PCG (b200spai_dist_t).  Collective: every rank constructs it and calls ``solve`` together.

    ``A``, ``G`` are the GLOBAL matrices resident on this rank's GPU (v1: every rank assembles the whole system and
    keeps only its rows); ``row_starts`` comes from ``partition_rows`` and must be identical on every rank."""

    def __init__(self, A: DeviceCSR, G: Optional[DeviceCSR], epsilon: float, row_starts: Sequence[int],
                 ext: Tuple[int, int], group=None, rank: Optional[int] = None, world: Optional[int] = None, mode: int = 0):
        self._lib = _lib.load()
        _lib.require_cuda()
        if world is None:
            import torch.distributed as dist

            world, rank = dist.get_world_size(group), dist.get_rank(group)
        self.rank, self.world = int(rank), int(world)
        n = A.shape[0]
        assert len(row_starts) == self.world + 1 and row_starts[0] == 0 and row_starts[-1] == n
        rs, re = int(row_starts[self.rank]), int(row_starts[self.rank + 1])
        self.layout = rank_layout(rs, re, int(ext[0]), int(ext[1]))
        self.layouts: List[Layout] = ([self.layout] if self.world == 1 else
                                      [tuple(int(v) for v in t) for t in _all_gather_object(self.layout, group)])
        ext_cap = max(l[3] - l[2] for l in self.layouts)
        self._h = C.c_void_p()
        _lib.check(self._lib.b200spai_dist_create(self.rank, self.world, ext_cap, C.byref(self._h)), "dist_create")
        handles = None
        if self.world > 1:
            hb = (C.c_ubyte * 64)()
            _lib.check(self._lib.b200spai_dist_ipc_handle(self._h, hb), "dist_ipc_handle")
            handles = b"".join(_all_gather_object(bytes(hb), group))
        lay = (C.c_int64 * (4 * self.world))(*[v for l in self.layouts for v in l])
        _lib.check(self._lib.b200spai_dist_connect(self._h, handles, lay), "dist_connect")
        # local matrices: owned rows, columns shifted into the extended layout
        _, _, shift, hi = self.layout
        self.n_local, self.ext_len = re - rs, hi - shift
        self.A = self._slice(A, rs, re, shift)
        self.G = self.GT = None
        if G is not None:
            GT = G.transpose()
            self.G = self._slice(G, rs, re, shift)
            self.GT = self._slice(GT, rs, re, shift)
            GT.close()
        _lib.check(self._lib.b200spai_dist_pcg_setup(self._h, self.A._h, self.G._h if self.G else None,
                                                     self.GT._h if self.GT else None, float(epsilon)), "dist_pcg_setup")
        if mode:
            _lib.check(self._lib.b200spai_dist_pcg_set_mode(self._h, mode), "dist_pcg_set_mode")
        self.group = group

    def _slice(self, M: DeviceCSR, rs: int, re: int, shift: int) -> DeviceCSR:
        h = C.c_void_p()
        _lib.check(self._lib.b200spai_csr_slice_rows(M._h, rs, re, shift, self.ext_len, _stream(), C.byref(h)), "csr_slice_rows")
        return DeviceCSR(h)

    def info(self):
        a, b, c, d = C.c_int32(), C.c_int32(), C.c_int64(), C.c_int32()
        _lib.check(self._lib.b200spai_dist_info(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)), "dist_info")
        return {"n_push": a.value, "n_recv": b.value, "push_elems": c.value, "owned_offset": d.value}

    def solve(self, b_local: torch.Tensor, precond: str = "ext_spai", rtol: float = 1e-6, max_iter: int = 0,
              history: bool = False):
        """b_local: the owned rows of the right-hand side.  Returns (x_local, PcgResult, history or None)."""
        assert b_local.is_cuda and b_local.dtype == torch.float64 and b_local.numel() == self.n_local
        if max_iter <= 0:
            max_iter = self.layouts[-1][1]  # n (validate.py:100-101)
        x = torch.empty(max(self.n_local, 1), dtype=torch.float64, device=b_local.device)[: self.n_local]
        hist = torch.zeros(max_iter + 1, dtype=torch.float64, device=b_local.device) if history else None
        res = _lib.PcgResult()
        _lib.check(self._lib.b200spai_dist_pcg_solve(self._h, PRECOND[precond], _ptr(b_local.contiguous()), _ptr(x), float(rtol),
                                                     int(max_iter), _ptr(hist), _stream(), C.byref(res)), "dist_pcg_solve")
        if hist is not None:
            hist = hist[: res.iterations + 1]
        return x, res, hist

    def gather(self, x_local: torch.Tensor) -> torch.Tensor:
        """The full solution vector on every rank (plumbing for checks; NCCL all_gather of ragged pieces)."""
        if self.world == 1:
            return x_local
        import torch.distributed as dist

        sizes = [l[1] - l[0] for l in self.layouts]
        pad = max(sizes)
        buf = torch.zeros(pad, dtype=x_local.dtype, device=x_local.device)
        buf[: x_local.numel()] = x_local
        out = [torch.empty_like(buf) for _ in sizes]
        dist.all_gather(out, buf, group=self.group)
        return torch.cat([o[:s] for o, s in zip(out, sizes)])

    def close(self):
        if self._h:
            self._lib.b200spai_dist_destroy(self._h)
            self._h = C.c_void_p()
        for m in (self.A, self.G, self.GT):
            if m is not None:
                m.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def row_weights_from_edges(edge_index: torch.Tensor, num_rows: int) -> torch.Tensor:
    """nnz_A + nnz_G + nnz_GT per row for pattern(G) = pattern(A) = the edge list: 2 * row degree + column degree."""
    row_deg = torch.bincount(edge_index[0], minlength=num_rows)
    col_deg = torch.bincount(edge_index[1], minlength=num_rows)
    return 2 * row_deg + col_deg
