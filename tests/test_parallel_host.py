"""Host-side logic of the row-partitioned solve (no GPU): the partitioner against a numpy mirror, the halo plan's
consistency between ranks, and -- over a real 2-process gloo group -- a distributed SpMV / PCG built from the same
layouts and plans, with scipy as the local operator, against the global scipy result.  The device path itself
(peer stores, flags) is covered by tests/test_dist_gpu.py on GPUs."""
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from learningsparsepreconditioner4gpu_b200 import parallel as par  # noqa: E402


def _grid_laplacian(n):
    I = sp.identity(n)
    T = sp.diags([-1.0, 2.0, -1.0], [-1, 0, 1], shape=(n, n))
    A = sp.kron(sp.kron(T, I), I) + sp.kron(sp.kron(I, T), I) + sp.kron(sp.kron(I, I), T)
    A = sp.csr_matrix(A + 0.01 * sp.identity(n ** 3))
    A.sort_indices()
    return A


def _edges(A):
    coo = A.tocoo()
    order = np.lexsort((coo.col, coo.row))
    return torch.from_numpy(np.stack([coo.row[order], coo.col[order]]).astype(np.int64))


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_partition_matches_numpy_mirror(world):
    rng = np.random.default_rng(world)
    w = rng.integers(0, 40, size=5000)
    starts = par.partition_rows(torch.from_numpy(w), world)
    c = np.cumsum(w)
    ref = [0] + [int(np.searchsorted(c, (k * int(c[-1])) // world, side="right")) for k in range(1, world)] + [len(w)]
    assert starts == ref
    assert starts == sorted(starts) and starts[0] == 0 and starts[-1] == len(w)
    loads = [int(w[a:b].sum()) for a, b in zip(starts[:-1], starts[1:])]
    assert max(loads) - min(loads) <= 2 * int(w.max())  # balanced up to one row either side


def test_partition_degenerate():
    z = par.partition_rows(torch.zeros(7, dtype=torch.int64), 3)  # zero weights: any monotone cut that covers the rows
    assert z[0] == 0 and z[-1] == 7 and z == sorted(z)
    s = par.partition_rows(torch.ones(2, dtype=torch.int64), 4)
    assert s[0] == 0 and s[-1] == 2 and s == sorted(s)


@pytest.mark.parametrize("world", [2, 4, 7])
def test_halo_plan_is_consistent_between_ranks(world):
    A = _grid_laplacian(9)
    ei = _edges(A)
    n = A.shape[0]
    starts = par.partition_rows(par.row_weights_from_edges(ei, n), world)
    layouts = [par.rank_layout(starts[r], starts[r + 1], *par.extended_range(ei, starts[r], starts[r + 1])) for r in range(world)]
    for r, (rs, re, shift, hi) in enumerate(layouts):
        assert (rs - shift) % 2 == 0 and shift <= rs and hi >= re
        # every column a rank's rows touch lies in its extended range
        cols = A[rs:re].indices
        assert cols.size == 0 or (cols.min() >= shift and cols.max() < hi)
    plans = [par.halo_plan(layouts, r) for r in range(world)]
    for r in range(world):
        pushes, recvs = plans[r]
        for (q, src_off, dst_off, length) in pushes:
            assert r in plans[q][1]                      # q expects data from r
            g0 = src_off + layouts[r][2]                 # global index of the first pushed entry
            assert layouts[r][0] <= g0 and g0 + length <= layouts[r][1]          # r owns it
            assert dst_off == g0 - layouts[q][2] and dst_off + length <= layouts[q][3] - layouts[q][2]
        for q in recvs:
            assert any(p[0] == r for p in plans[q][0])   # ... and q pushes to r
    # the pushes cover every non-owned column some rank reads
    for r, (rs, re, shift, hi) in enumerate(layouts):
        have = np.zeros(hi - shift, dtype=bool)
        have[rs - shift: re - shift] = True
        for q in range(world):
            for (dst, src_off, dst_off, length) in plans[q][0]:
                if dst == r:
                    have[dst_off: dst_off + length] = True
        cols = np.unique(A[rs:re].indices) - shift
        assert have[cols].all()


def _worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = _grid_laplacian(8)
        ei = _edges(A)
        n = A.shape[0]
        starts = par.partition_rows(par.row_weights_from_edges(ei, n), world)
        mine = par.rank_layout(starts[rank], starts[rank + 1], *par.extended_range(ei, starts[rank], starts[rank + 1]))
        layouts = par._all_gather_object(mine, None)
        rs, re, shift, hi = mine
        pushes, recvs = par.halo_plan(layouts, rank)
        A_loc = sp.csr_matrix(A[rs:re, shift if shift >= 0 else 0: hi])
        pad = 0 if shift >= 0 else -shift                 # shift may be -1: one unused slot in front
        ext = np.zeros(hi - shift)

        def exchange(owned):
            """emulates dist_halo_kernel with gloo point-to-point: push my slices, receive the neighbours'."""
            ext[rs - shift: re - shift] = owned
            reqs = []
            for (qr, src_off, dst_off, length) in pushes:
                reqs.append(dist.isend(torch.from_numpy(ext[src_off: src_off + length].copy()), qr))
            for qr in recvs:
                (_, s_off, d_off, length) = next(p for p in par.halo_plan(layouts, qr)[0] if p[0] == rank)
                buf = torch.empty(length, dtype=torch.float64)
                dist.recv(buf, qr)
                ext[d_off: d_off + length] = buf.numpy()
            for r_ in reqs:
                r_.wait()
            return ext[pad:]

        def allreduce(v):
            """emulates dist_reduce_kernel: every rank sums the per-rank values in rank order."""
            vals = par._all_gather_object(float(v), None)
            t = 0.0
            for x in vals:
                t += x
            return t

        rng = np.random.default_rng(0)
        xg = rng.standard_normal(n)
        y_loc = A_loc @ exchange(xg[rs:re])
        ok_spmv = np.allclose(y_loc, (A @ xg)[rs:re], rtol=0, atol=1e-12)
        # unpreconditioned CG with the scipy recurrence, distributed
        b = (A @ np.ones(n))[rs:re]
        x = np.zeros(re - rs)
        r = b.copy()
        bn = np.sqrt(allreduce(r @ r))
        it, rho_prev, p = 0, 0.0, None
        while np.sqrt(allreduce(r @ r)) >= 1e-8 * bn and it < 500:
            rho = allreduce(r @ r)
            p = r.copy() if it == 0 else r + (rho / rho_prev) * p
            qv = A_loc @ exchange(p)
            alpha = rho / allreduce(p @ qv)
            x += alpha * p
            r -= alpha * qv
            rho_prev = rho
            it += 1
        err = np.abs(x - 1.0).max()
        q.put((rank, bool(ok_spmv), it, float(err)))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_distributed_spmv_and_cg():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _, _ in res)
    assert res[0][2] == res[1][2] and res[0][2] < 500          # same iteration count on both ranks
    assert max(e for *_, e in res) < 1e-6


# ---------------------------------------------------------------------------------------------------------
# partitioned GNN / assembly: edge ownership, G boundary swap, node-feature halo refresh
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("world", [1, 2, 3, 5])
def test_g_exchange_plan_covers_every_edge_once(world):
    A = _grid_laplacian(6)
    ei = _edges(A)
    n, E = A.shape[0], ei.shape[1]
    starts = par.partition_rows(par.row_weights_from_edges(ei, n), world)
    plans = [par.g_exchange_plan(ei, starts, r) for r in range(world)]
    G = torch.arange(E, dtype=torch.float64)                  # value = global edge id
    seen_t = torch.zeros(E, dtype=torch.int64)
    for r, (e0, e1, ids_t, loc_t, loc_r, send, recv) in enumerate(plans):
        rs, re = starts[r], starts[r + 1]
        assert bool(((ei[0][e0:e1] >= rs) & (ei[0][e0:e1] < re)).all()) and e1 - e0 == int(((ei[0] >= rs) & (ei[0] < re)).sum())
        assert bool(((ei[1][ids_t] >= rs) & (ei[1][ids_t] < re)).all())
        seen_t[ids_t] += 1
        # the row slice is filled exactly once: locally or by one peer, with the right global edge
        out = torch.full((e1 - e0,), -1.0, dtype=torch.float64)
        out[loc_r] = G[ids_t][loc_t]
        for q, idx in recv.items():
            assert bool((out[idx] == -1).all())
            out[idx] = G[plans[q][2]][plans[q][5][r]]           # what q sends to r, in q's order
        assert torch.equal(out, G[e0:e1])
        for q, idx in send.items():
            assert r in plans[q][6] and plans[q][6][r].numel() == idx.numel()
    assert bool((seen_t == 1).all())                          # every edge is computed by exactly one rank


def test_halo_ranges_mirror_halo_plan():
    A = _grid_laplacian(7)
    ei = _edges(A)
    n = A.shape[0]
    for world in (2, 4):
        starts = par.partition_rows(par.row_weights_from_edges(ei, n), world)
        layouts = [par.rank_layout(starts[r], starts[r + 1], *par.extended_range(ei, starts[r], starts[r + 1])) for r in range(world)]
        for r in range(world):
            sends, recvs = par.halo_ranges(layouts, r)
            pushes, precv = par.halo_plan(layouts, r)
            assert [(q, o, l) for q, o, _, l in pushes] == sends and [q for q, _, _ in recvs] == precv
            for q, off, length in recvs:                      # the matching send of q has the same length
                (m,) = [s for s in par.halo_ranges(layouts, q)[0] if s[0] == r]
                assert m[2] == length and m[1] + layouts[q][2] == off + layouts[r][2]


def _worker_exchange(rank, world, port, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        A = _grid_laplacian(7)
        ei = _edges(A)
        n, E = A.shape[0], ei.shape[1]
        starts = par.partition_rows(par.row_weights_from_edges(ei, n), world)
        gen = torch.Generator().manual_seed(5)
        G = torch.randn(E, 3, 3, generator=gen)               # same "global G" on every rank, block size 3
        X = torch.randn(n, 16, generator=gen)                 # same global node features
        plan = par.g_exchange_plan(ei, starts, rank)
        e0, e1, ids_t = plan[:3]
        rows = par.exchange_g_rows(G[ids_t].contiguous(), plan)
        ok_g = torch.equal(rows, G[e0:e1])
        mine = par.rank_layout(starts[rank], starts[rank + 1], *par.extended_range(ei, starts[rank], starts[rank + 1]))
        layouts = [tuple(t) for t in par._all_gather_object(mine, None)]
        rs, re, shift, hi = mine
        lo = max(shift, 0)
        buf = torch.full((hi - shift, 16), float("nan"))
        buf[rs - shift: re - shift] = X[rs:re]                # only the owned rows are valid
        par.exchange_halo_rows(buf, layouts, rank)
        ok_x = torch.equal(buf[lo - shift:], X[lo:hi])
        q.put((rank, bool(ok_g), bool(ok_x)))
    finally:
        dist.destroy_process_group()


def test_gloo_world3_g_swap_and_feature_halo():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker_exchange, args=(r, 3, port, q)) for r in range(3)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(a and b for _, a, b in res), res


def _random_spd_pattern(n, density, seed):
    rng = np.random.default_rng(seed)
    M = sp.random(n, n, density=density, random_state=rng, format="csr")
    A = sp.csr_matrix(M.T @ M + sp.identity(n))
    A.sort_indices()
    return A


@pytest.mark.parametrize("world", [1, 3, 4])
@pytest.mark.parametrize("kind", ["grid", "irregular"])
def test_partitioned_system_local_edge_lists(kind, world):
    """PartitionedSystem's index side on CPU (no kernels): the row-owned list reproduces the owned rows of the matrix in
    the extended column window, the target-owned list reproduces the owned rows of the TRANSPOSE (what G^T is assembled
    from), both are row-major / column-ascending as b200spai_edges_to_csr requires, and ext_slice pads a negative shift."""
    A = _grid_laplacian(6) if kind == "grid" else _random_spd_pattern(300, 0.01, 7)
    ei = _edges(A)
    n, E = A.shape[0], ei.shape[1]
    vals = torch.arange(1, E + 1, dtype=torch.float64)        # distinct value per edge
    Aval = sp.csr_matrix((vals.numpy(), (ei[0].numpy(), ei[1].numpy())), shape=(n, n))
    starts = par.partition_rows(par.row_weights_from_edges(ei, n), world)
    for rank in range(world):
        ps = par.PartitionedSystem.__new__(par.PartitionedSystem)
        # the constructor's collective (all_gather of layouts) needs a process group: build the layouts by hand
        lay = [par.rank_layout(starts[r], starts[r + 1], *par.extended_range(ei, starts[r], starts[r + 1])) for r in range(world)]
        ps.rank, ps.world, ps.group = rank, world, None
        ps.N, ps.b, ps.starts = n, 1, list(starts)
        ps.layout_nodes, ps.layouts_nodes = lay[rank], lay
        rs, re, shift, hi = lay[rank]
        ps.shift, ps.hi, ps.ext_nodes = shift, hi, hi - shift
        ps.plan = par.g_exchange_plan(ei, starts, rank)
        ps.e0, ps.e1, ps.ids_t = ps.plan[0], ps.plan[1], ps.plan[2]
        ei_t = ei[:, ps.ids_t] - shift
        ei_r = ei[:, ps.e0:ps.e1] - shift
        order = torch.argsort(ei_t[1], stable=True)
        ei_tT = torch.stack([ei_t[1][order], ei_t[0][order]])
        ext = hi - shift
        for e in (ei_r, ei_tT):                               # sorted row-major / column-ascending, in range
            if e.shape[1]:
                key = e[0] * ext + e[1]
                assert bool((key[1:] > key[:-1]).all()) and int(e.min()) >= 0 and int(e.max()) < ext
        lo = max(shift, 0)
        # owned rows of A from the row-owned list
        L = sp.csr_matrix((vals[ps.e0:ps.e1].numpy(), (ei_r[0].numpy(), ei_r[1].numpy())), shape=(ext, ext))[rs - shift:re - shift]
        ref = sp.hstack([sp.csr_matrix((re - rs, lo - shift)), Aval[rs:re, lo:hi]]).tocsr()
        assert (abs(L - ref)).nnz == 0 and L.nnz == Aval[rs:re].nnz
        # owned rows of A^T from the target-owned list (values permuted with the same stable order)
        vt = vals[ps.ids_t][order]
        LT = sp.csr_matrix((vt.numpy(), (ei_tT[0].numpy(), ei_tT[1].numpy())), shape=(ext, ext))[rs - shift:re - shift]
        refT = sp.hstack([sp.csr_matrix((re - rs, lo - shift)), Aval.T.tocsr()[rs:re, lo:hi]]).tocsr()
        assert (abs(LT - refT)).nnz == 0
        # ext_slice: rows [shift, hi) with zero rows in front when shift < 0
        x = torch.arange(n * 2, dtype=torch.float32).reshape(n, 2) + 1
        xs = ps.ext_slice(x)
        assert xs.shape == (ext, 2) and torch.equal(xs[lo - shift:], x[lo:hi]) and float(xs[: lo - shift].abs().sum()) == 0.0
        assert ps.row_starts == list(starts) and ps.layout == lay[rank]


@pytest.mark.parametrize("world", [1, 3])
def test_partitioned_system_from_rank_slices_equals_global(world, monkeypatch):
    """The constructor fed with this rank's two pieces of the edge list (``rank_edge_slices``: what a distributed caller
    hands each rank) builds the same layout, local edge lists and G-exchange plan as the one fed with the global list."""
    A = _random_spd_pattern(400, 0.012, 3)
    ei = _edges(A)
    n = A.shape[0]
    starts = par.partition_rows(par.row_weights_from_edges(ei, n), world)
    lay = [par.rank_layout(starts[r], starts[r + 1], *par.extended_range(ei, starts[r], starts[r + 1])) for r in range(world)]
    monkeypatch.setattr(par, "_all_gather_object", lambda obj, group: list(lay))
    for rank in range(world):
        g = par.PartitionedSystem(ei, n, 1, starts, rank=rank, world=world)
        e0, e1, ids_t = par.rank_edge_slices(ei, starts, rank)
        s = par.PartitionedSystem(None, n, 1, starts, rank=rank, world=world, slices=(ei[:, e0:e1].clone(), ei[:, ids_t].clone()))
        assert g.layout_nodes == s.layout_nodes == lay[rank] and g.ext_nodes == s.ext_nodes
        for name in ("ei_t", "ei_r", "ei_tT", "order_t"):
            assert torch.equal(getattr(g, name), getattr(s, name)), name
        assert torch.equal(g.plan[3], s.plan[3]) and torch.equal(g.plan[4], s.plan[4])
        assert set(g.plan[5]) == set(s.plan[5]) and set(g.plan[6]) == set(s.plan[6])
        for q in g.plan[5]:
            assert torch.equal(g.plan[5][q], s.plan[5][q])
        for q in g.plan[6]:
            assert torch.equal(g.plan[6][q], s.plan[6][q])
        ref = par.g_exchange_plan(ei, starts, rank)            # and both equal the stand-alone plan function
        assert torch.equal(g.plan[3], ref[3]) and torch.equal(g.plan[4], ref[4])
        assert (s.n_edges_r, s.n_edges_t) == (e1 - e0, int(ids_t.numel()))
