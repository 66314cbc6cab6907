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
import os
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

    _connections: dict = {}   # (rank, world, group id) -> (handle, capacity): parked by close() when reuse=True

    @classmethod
    def drop_connections(cls):
        """Tear down the parked connections (collective: call on every rank, after a barrier)."""
        lib = _lib.load()
        for h, _ in cls._connections.values():
            lib.b200spai_dist_destroy(h)
        cls._connections.clear()

    def __init__(self, A: Optional[DeviceCSR], G: Optional[DeviceCSR], epsilon: float, row_starts: Sequence[int],
                 ext: Tuple[int, int], group=None, rank: Optional[int] = None, world: Optional[int] = None, mode: int = 0,
                 local: Optional[Tuple[DeviceCSR, Optional[DeviceCSR], Optional[DeviceCSR]]] = None,
                 layout: Optional[Layout] = None, reuse: bool = False):
        """``local`` = (A_loc, G_loc, GT_loc): the owned rows already assembled in the extended column layout
        (``PartitionedSystem.assemble``); then ``A`` / ``G`` are ignored and nothing global is needed on this rank.
        ``layout`` overrides the layout derived from ``ext`` (block systems keep it node-aligned).
        ``reuse``: ``close`` parks the connection (peer-visible segment + IPC mappings, ~70 ms to set up) for the next
        ``DistPCG(..., reuse=True)`` of the same ranks instead of tearing it down -- every rank must pass the same flag."""
        self._lib = _lib.load()
        _lib.require_cuda()
        if world is None:
            import torch.distributed as dist

            world, rank = dist.get_world_size(group), dist.get_rank(group)
        self.rank, self.world = int(rank), int(world)
        n = int(row_starts[-1]) if local is not None else A.shape[0]
        assert len(row_starts) == self.world + 1 and row_starts[0] == 0 and row_starts[-1] == n
        rs, re = int(row_starts[self.rank]), int(row_starts[self.rank + 1])
        self.layout = tuple(int(v) for v in layout) if layout is not None else rank_layout(rs, re, int(ext[0]), int(ext[1]))
        assert self.layout[0] == rs and self.layout[1] == re
        self.layouts: List[Layout] = ([self.layout] if self.world == 1 else
                                      [tuple(int(v) for v in t) for t in _all_gather_object(self.layout, group)])
        ext_cap = max(l[3] - l[2] for l in self.layouts)
        lay = (C.c_int64 * (4 * self.world))(*[v for l in self.layouts for v in l])
        self._key = (self.rank, self.world, id(group))
        cached = DistPCG._connections.pop(self._key, None) if reuse else None
        if cached is not None and cached[1] < ext_cap:      # identical decision on every rank (ext_cap is global)
            self._lib.b200spai_dist_destroy(cached[0])
            cached = None
        self._reuse = bool(reuse)
        if cached is not None:
            # the CUDA-IPC mappings of the peers' segments survive from the previous system: only the plan changes
            self._h, self._cap = cached
            _lib.check(self._lib.b200spai_dist_relayout(self._h, lay), "dist_relayout")
        else:
            self._cap = ext_cap + ext_cap // 4 if reuse else ext_cap
            self._h = C.c_void_p()
            _lib.check(self._lib.b200spai_dist_create(self.rank, self.world, self._cap, C.byref(self._h)), "dist_create")
            handles = None
            if self.world > 1:
                hb = (C.c_ubyte * 64)()
                _lib.check(self._lib.b200spai_dist_ipc_handle(self._h, hb), "dist_ipc_handle")
                handles = b"".join(_all_gather_object(bytes(hb), group))
            _lib.check(self._lib.b200spai_dist_connect(self._h, handles, lay), "dist_connect")
        # local matrices: owned rows, columns shifted into the extended layout
        _, _, shift, hi = self.layout
        self.n_local, self.ext_len = re - rs, hi - shift
        self.G = self.GT = None
        if local is not None:
            self.A, self.G, self.GT = local
            for m in local:
                assert m is None or m.shape == (self.n_local, self.ext_len), (m.shape, self.n_local, self.ext_len)
        else:
            self.A = self._slice(A, rs, re, shift)
        if local is None and G is not None:
            GT = G.transpose()
            self.G = self._slice(G, rs, re, shift)
            self.GT = self._slice(GT, rs, re, shift)
            GT.close()
        _lib.check(self._lib.b200spai_dist_pcg_setup(self._h, self.A._h, self.G._h if self.G else None,
                                                     self.GT._h if self.GT else None, float(epsilon)), "dist_pcg_setup")
        # mode 4 (one persistent kernel per rank, halo + all-reduce inside its grid barriers) must be taken by ALL ranks
        # or by none: every rank asks the library whether its own matrices qualify and the answers are and-ed
        if self.G is not None and mode in (0, 4) and not os.environ.get("B200SPAI_NO_FUSED"):
            ok = C.c_int32()
            _lib.check(self._lib.b200spai_dist_pcg_fused_ok(self._h, _stream(), C.byref(ok)), "dist_pcg_fused_ok")
            oks = [ok.value] if self.world == 1 else _all_gather_object(int(ok.value), group)
            if all(v == 2 for v in oks) or (mode == 4 and all(oks)):   # auto: only in the streaming regime (large systems)
                mode = 4
            elif mode == 4:
                raise _lib.B200SpaiError(f"DistPCG: mode 4 requested but ranks {[i for i, v in enumerate(oks) if not v]} cannot run it")
        elif mode == 4:
            raise _lib.B200SpaiError("DistPCG: mode 4 needs G (ext_spai) and B200SPAI_NO_FUSED unset")
        self.mode = mode
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

    PROFILE_KEYS = ("GT_r", "halo_t", "G_t", "allreduce_rho+push_z", "update_p", "A_p", "allreduce_pq+push_q", "update_xr")

    def profile_iterations(self, b_local: torch.Tensor, iterations: int = 50, precond: str = "ext_spai") -> dict:
        """Mean device time (ms) of every launch of one iteration (collective; measurement only)."""
        x = torch.empty(max(self.n_local, 1), dtype=torch.float64, device=b_local.device)
        out = (C.c_float * 8)()
        _lib.check(self._lib.b200spai_dist_pcg_profile_iterations(self._h, PRECOND[precond], _ptr(b_local.contiguous()), _ptr(x),
                                                                  int(iterations), out, _stream()), "dist_pcg_profile_iterations")
        return {k: round(float(v), 5) for k, v in zip(self.PROFILE_KEYS, out)}

    def fused_profile(self):
        """Mode 4, after a solve: ([ms in t = G^T r, z = G t, p update, q = A p, x/r update] incl. the barrier closing
        each phase, iterations covered), this rank's view."""
        ms = (C.c_double * 5)()
        it = C.c_int32()
        _lib.check(self._lib.b200spai_dist_pcg_fused_profile(self._h, ms, C.byref(it), _stream()), "dist_pcg_fused_profile")
        return [float(v) for v in ms], it.value

    def fused_barrier_profile(self):
        """Mode 4, world > 1, after a solve: {mean us CTA 0 waited for its own grid, mean us from there to the last peer}
        per cross-rank barrier, and the number of barriers."""
        us = (C.c_double * 2)()
        nb = C.c_int64()
        _lib.check(self._lib.b200spai_dist_pcg_fused_barrier_profile(self._h, us, C.byref(nb), _stream()), "dist_pcg_fused_barrier_profile")
        return {"wait_local_grid_us": round(us[0], 3), "wait_peers_us": round(us[1], 3), "barriers": nb.value}

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
            if getattr(self, "_reuse", False):
                old = DistPCG._connections.pop(self._key, None)
                if old is not None:
                    self._lib.b200spai_dist_destroy(old[0])
                DistPCG._connections[self._key] = (self._h, self._cap)
            else:
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


# ---------------------------------------------------------------------------------------------------------
# partitioned GNN + assembly: every rank computes only its share of G and of the matrices
# ---------------------------------------------------------------------------------------------------------
def owner_of(index: torch.Tensor, starts: Sequence[int]) -> torch.Tensor:
    """Rank that owns each (node / row) index under the contiguous partition ``starts``."""
    bounds = torch.tensor(list(starts[1:-1]), dtype=torch.int64, device=index.device)
    return torch.bucketize(index, bounds, right=True)


def owned_edge_sets(edge_index: torch.Tensor, rs: int, re: int):
    """(e0, e1, ids_t): the edges whose ROW is owned are the contiguous slice [e0, e1) of the row-major edge list (they
    give the owned rows of A and G); ``ids_t`` = ascending ids of the edges whose TARGET ``edge_index[1]`` is owned --
    the edges this rank runs the GNN on, because messages are summed at the target (basic_layers.py:193-225)."""
    row, col = edge_index[0], edge_index[1]
    bnd = torch.tensor([rs, re], dtype=row.dtype, device=row.device)
    e0, e1 = (int(v) for v in torch.searchsorted(row.contiguous(), bnd).tolist())
    ids_t = torch.nonzero((col >= rs) & (col < re)).reshape(-1)
    return e0, e1, ids_t


def g_exchange_plan(edge_index: torch.Tensor, starts: Sequence[int], rank: int):
    """Who needs which G values.  The GNN leaves rank r with G on its target-owned edges (columns of G), the solver
    needs the owned ROWS of G: entry (i, j) with i owned by q and j owned by r != q travels r -> q.

    Returns ``(e0, e1, ids_t, loc_t, loc_r, send, recv)``: ``loc_t`` / ``loc_r`` = positions (in ``ids_t`` / in the
    row slice) of the edges that are both; ``send[q]`` = positions in ``ids_t`` to ship to q; ``recv[q]`` = positions in
    the row slice filled by q.  Both sides enumerate a (row owner, column owner) class in ascending edge id, so the
    k-th value sent is the k-th value expected."""
    rs, re = int(starts[rank]), int(starts[rank + 1])
    e0, e1, ids_t = owned_edge_sets(edge_index, rs, re)
    row_owner_t = owner_of(edge_index[0][ids_t], starts)
    col_owner_r = owner_of(edge_index[1][e0:e1], starts)
    loc_t = torch.nonzero(row_owner_t == rank).reshape(-1)
    loc_r = torch.nonzero(col_owner_r == rank).reshape(-1)
    send, recv = {}, {}
    for q in range(len(starts) - 1):
        if q == rank:
            continue
        s = torch.nonzero(row_owner_t == q).reshape(-1)
        if s.numel():
            send[q] = s
        r = torch.nonzero(col_owner_r == q).reshape(-1)
        if r.numel():
            recv[q] = r
    return e0, e1, ids_t, loc_t, loc_r, send, recv


def halo_ranges(layouts: Sequence[Layout], rank: int):
    """(sends, recvs) as [(peer, offset, length)] in this rank's extended index space: the owned slices a peer's
    extended range contains, and the slices of this rank's halo each peer owns (``halo_plan`` seen from both ends)."""
    rs, re, shift, hi = layouts[rank]
    sends, recvs = [], []
    for q, (rs_q, re_q, lo_q, hi_q) in enumerate(layouts):
        if q == rank:
            continue
        a, b = max(rs, lo_q), min(re, hi_q)
        if a < b:
            sends.append((q, a - shift, b - a))
        a2, b2 = max(rs_q, shift), min(re_q, hi)
        if a2 < b2:
            recvs.append((q, a2 - shift, b2 - a2))
    return sends, recvs


def _p2p(sends, recvs, group):
    """sends = [(peer, tensor)], recvs = [(peer, tensor)] (contiguous); one batched NCCL / gloo exchange."""
    import torch.distributed as dist

    ops = [dist.P2POp(dist.isend, t, q, group) for q, t in sends] + [dist.P2POp(dist.irecv, t, q, group) for q, t in recvs]
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()


def exchange_halo_rows(buf: torch.Tensor, layouts: Sequence[Layout], rank: int, group=None) -> None:
    """Refresh the halo rows of ``buf`` ([extended length, ...], contiguous) with their owners' rows."""
    sends, recvs = halo_ranges(layouts, rank)
    _p2p([(q, buf[o:o + l]) for q, o, l in sends], [(q, buf[o:o + l]) for q, o, l in recvs], group)


def exchange_g_rows(g_t: torch.Tensor, plan, group=None) -> torch.Tensor:
    """G values [E_t, ...] on the target-owned edges -> G values on the row-owned edges [e1 - e0, ...]."""
    e0, e1, ids_t, loc_t, loc_r, send, recv = plan
    out = torch.empty((e1 - e0,) + tuple(g_t.shape[1:]), dtype=g_t.dtype, device=g_t.device)
    out[loc_r] = g_t[loc_t]
    sb = [(q, g_t[idx].contiguous()) for q, idx in sorted(send.items())]
    rb = [(q, torch.empty((idx.numel(),) + tuple(g_t.shape[1:]), dtype=g_t.dtype, device=g_t.device))
          for q, idx in sorted(recv.items())]
    _p2p(sb, rb, group)
    for q, t in rb:
        out[recv[q]] = t
    return out


def rank_edge_slices(edge_index: torch.Tensor, starts: Sequence[int], rank: int):
    """(e0, e1, ids_t) of ``owned_edge_sets`` for ``rank``: the input partitioner.  A distributed caller runs it once per
    topology where the global edge list lives and hands each rank ``edge_index[:, e0:e1]``, ``edge_index[:, ids_t]``,
    ``matrix_values[e0:e1]``, ``edge_attr[ids_t]`` and the node window of ``x`` / ``mask``."""
    return owned_edge_sets(edge_index, int(starts[rank]), int(starts[rank + 1]))


class PartitionedSystem:
    """One rank's share of a row-partitioned system: the GNN runs on the edges whose target node this
    rank owns (node-feature halo refreshed after every MP layer), the boundary G values are swapped once, and A, G, G^T
    are assembled for the owned rows only -- nothing of size E or N is computed twice.

    ``edge_index`` is the global [2, E] list (row-major, column-ascending) on this rank's GPU; ``starts`` the NODE
    partition from ``partition_rows``.  Collective: every rank of ``group`` builds it and calls the methods together."""

    def __init__(self, edge_index: Optional[torch.Tensor], num_nodes: int, block_size: int, starts: Sequence[int],
                 rank: Optional[int] = None, world: Optional[int] = None, group=None,
                 slices: Optional[Tuple[torch.Tensor, torch.Tensor]] = None):
        """``slices`` = (ei_r, ei_t): this rank's two pieces of the GLOBAL edge list, in global node numbering and ascending
        edge order -- the edges whose row it owns (a contiguous run of the row-major list) and the edges whose target
        (column) it owns (``rank_edge_slices``).  With ``slices`` nothing of global size E is needed on this rank:
        ``edge_index`` may be None and the ``*_local`` methods take per-rank inputs."""
        if world is None:
            import torch.distributed as dist

            world, rank = dist.get_world_size(group), dist.get_rank(group)
        self.rank, self.world, self.group = int(rank), int(world), group
        self.N, self.b, self.starts = int(num_nodes), int(block_size), [int(v) for v in starts]
        assert len(self.starts) == self.world + 1 and self.starts[0] == 0 and self.starts[-1] == self.N
        rs, re = self.starts[self.rank], self.starts[self.rank + 1]
        if slices is None:
            e0, e1, ids_t = owned_edge_sets(edge_index, rs, re)
            ei_r, ei_t = edge_index[:, e0:e1], edge_index[:, ids_t]
            self.e0, self.e1, self.ids_t = e0, e1, ids_t
        else:
            ei_r, ei_t = slices
            self.e0, self.e1, self.ids_t = None, None, None       # global positions are not known (and not needed)
        self.n_edges_r, self.n_edges_t = int(ei_r.shape[1]), int(ei_t.shape[1])
        # extended node window: owned rows + every column the owned rows touch + every row that touches an owned column
        lo, hi = rs, re
        if self.n_edges_r:
            lo, hi = min(lo, int(ei_r[1].min())), max(hi, int(ei_r[1].max()) + 1)
        if self.n_edges_t:
            lo, hi = min(lo, int(ei_t[0].min())), max(hi, int(ei_t[0].max()) + 1)
        self.layout_nodes = rank_layout(rs, re, lo, hi)
        self.layouts_nodes: List[Layout] = ([self.layout_nodes] if self.world == 1 else
                                            [tuple(int(v) for v in t) for t in _all_gather_object(self.layout_nodes, group)])
        self.shift, self.hi = self.layout_nodes[2], hi
        self.ext_nodes = hi - self.shift
        # who needs which G values (g_exchange_plan on the two slices)
        row_owner_t = owner_of(ei_t[0], self.starts)
        col_owner_r = owner_of(ei_r[1], self.starts)
        loc_t = torch.nonzero(row_owner_t == self.rank).reshape(-1)
        loc_r = torch.nonzero(col_owner_r == self.rank).reshape(-1)
        send, recv = {}, {}
        for q in range(self.world):
            if q == self.rank:
                continue
            sq = torch.nonzero(row_owner_t == q).reshape(-1)
            if sq.numel():
                send[q] = sq
            rq = torch.nonzero(col_owner_r == q).reshape(-1)
            if rq.numel():
                recv[q] = rq
        self.plan = (0, self.n_edges_r, None, loc_t, loc_r, send, recv)
        # local edge lists in extended node numbering (global - shift)
        self.ei_t = (ei_t - self.shift).contiguous()          # target-owned, for the GNN and G^T
        self.ei_r = (ei_r - self.shift).contiguous()          # row-owned, for A and G
        order = torch.argsort(self.ei_t[1], stable=True)      # by target, rows ascending inside
        self.order_t = order
        self.ei_tT = torch.stack([self.ei_t[1][order], self.ei_t[0][order]]).contiguous()
        self._graph = None

    # scalar-row quantities for the solver ------------------------------------------------------------------
    @property
    def row_starts(self) -> List[int]:
        return [v * self.b for v in self.starts]

    @property
    def layout(self) -> Layout:
        return tuple(v * self.b for v in self.layout_nodes)

    def ext_slice(self, t: torch.Tensor) -> torch.Tensor:
        """Rows [shift, hi) of a per-node tensor, zero-padded where shift < 0."""
        lo = max(self.shift, 0)
        body = t[lo:self.hi]
        if lo == self.shift:
            return body.contiguous()
        pad = torch.zeros((lo - self.shift,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        return torch.cat([pad, body]).contiguous()

    # GNN ---------------------------------------------------------------------------------------------------
    def gnn_forward(self, gnn, x: torch.Tensor, edge_attr: torch.Tensor) -> torch.Tensor:
        """G values [E_t, b*b] of the target-owned edges (``ids_t`` order) from the GLOBAL x / edge_attr; bit-identical
        to the single-GPU forward on those edges (per target the messages are summed in the same ascending edge order)."""
        assert self.ids_t is not None, "built from slices: use gnn_forward_local with this rank's inputs"
        return self.gnn_forward_local(gnn, self.ext_slice(x.detach().to(torch.float32)),
                                      edge_attr.detach().to(torch.float32)[self.ids_t])

    def gnn_forward_local(self, gnn, x_ext: torch.Tensor, ea_t: torch.Tensor) -> torch.Tensor:
        """The same from this rank's inputs only: ``x_ext`` = node features of the extended window [shift, hi) (zero rows
        where shift < 0, ``ext_slice``), ``ea_t`` = edge features of the target-owned edges."""
        from .neural_cg.nn.gnns import DeviceGraph, F

        lib = _lib.load()
        if self._graph is None:
            self._graph = DeviceGraph(self.ei_t, self.ext_nodes)
        g = self._graph
        x = x_ext
        x_loc = x_ext.detach().to(torch.float32).contiguous()
        ea_loc = ea_t.detach().to(torch.float32).contiguous()
        assert x_loc.shape[0] == self.ext_nodes and ea_loc.shape[0] == self.n_edges_t
        out = torch.empty((self.n_edges_t, gnn.edge_out), dtype=torch.float32, device=x.device)
        eng, L = gnn._get_engine(), gnn.num_mp_layers
        for stage in range(L + 2):
            _lib.check(lib.b200spai_gnn_forward_stage(eng, g._h, stage, _ptr(x_loc), _ptr(ea_loc), _ptr(out), _stream()),
                       "gnn_forward_stage")
            if 1 <= stage <= L and self.world > 1:
                p = C.c_void_p()
                _lib.check(lib.b200spai_gnn_node_features(g._h, stage, C.byref(p)), "gnn_node_features")
                exchange_halo_rows(_wrap_f32(p.value, (self.ext_nodes, F), x.device), self.layouts_nodes, self.rank, self.group)
        return out

    def g_rows(self, g_t: torch.Tensor) -> torch.Tensor:  # collective
        """G values of the row-owned edges (global edge order e0..e1) after the boundary swap."""
        return exchange_g_rows(g_t, self.plan, self.group)

    # assembly ----------------------------------------------------------------------------------------------
    def _local_rows(self, ei: torch.Tensor, vals: torch.Tensor, mask_ext: Optional[torch.Tensor]) -> DeviceCSR:
        """to_csr_cpu semantics on the extended node window, then the owned rows."""
        full = DeviceCSR.from_edges(ei, vals, self.ext_nodes, mask_ext)
        rs, re, shift, _ = self.layout
        h = C.c_void_p()
        _lib.check(_lib.load().b200spai_csr_slice_rows(full._h, rs - shift, re - shift, 0, self.ext_nodes * self.b, _stream(),
                                                       C.byref(h)), "csr_slice_rows")
        full.close()
        return DeviceCSR(h)

    def assemble(self, matrix_values: torch.Tensor, g_t: torch.Tensor, mask: Optional[torch.Tensor]):
        """(A_loc, G_loc, GT_loc): owned rows of the masked fp64 CSR matrices, columns in the extended layout.
        ``matrix_values`` [E, b, b] global (only rows e0..e1 are read), ``g_t`` from ``gnn_forward``."""
        assert self.e0 is not None, "built from slices: use assemble_local with this rank's inputs"
        b = self.b
        mask_ext = None if mask is None else self.ext_slice(mask.detach().to(torch.float32).reshape(self.N, b))
        return self.assemble_local(matrix_values.detach().reshape(-1, b, b)[self.e0:self.e1], g_t, mask_ext)

    def assemble_local(self, mv_r: torch.Tensor, g_t: torch.Tensor, mask_ext: Optional[torch.Tensor]):
        """The same from this rank's inputs only: ``mv_r`` [E_r, b, b] = matrix values of the row-owned edges,
        ``mask_ext`` [ext_nodes, b] = mask over the extended window."""
        b = self.b
        if mask_ext is not None:
            mask_ext = mask_ext.detach().to(torch.float32).reshape(self.ext_nodes, b).contiguous()
        a_vals = mv_r.detach().to(torch.float32).reshape(-1, b, b).contiguous()
        A_loc = self._local_rows(self.ei_r, a_vals, mask_ext)
        g_t = g_t.reshape(-1, b, b)
        G_loc = self._local_rows(self.ei_r, self.g_rows(g_t).contiguous(), mask_ext)
        GT_loc = self._local_rows(self.ei_tT, g_t[self.order_t].transpose(1, 2).contiguous(), mask_ext)
        return A_loc, G_loc, GT_loc

    def make_solver(self, local, epsilon: float, mode: int = 0, reuse: bool = False) -> "DistPCG":
        return DistPCG(None, None, epsilon, self.row_starts, (0, 0), group=self.group, rank=self.rank, world=self.world,
                       mode=mode, local=local, layout=self.layout, reuse=reuse)

    def close(self):
        if self._graph is not None:
            self._graph.close()
            self._graph = None


def _wrap_f32(ptr: int, shape, device) -> torch.Tensor:
    """Torch view of library-owned device memory (no copy) through the CUDA array interface."""
    import math

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 2,
                                  "strides": None}
    assert math.prod(shape) > 0
    return torch.as_tensor(h, device=device)
