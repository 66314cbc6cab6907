"""GPU tests of the row-partitioned solver (csrc/dist.cu) through the C ABI.

world = 1 exercises the whole device path (extended layout, all-reduce and halo kernels with no peers, graph and stream
drivers) on one GPU and must reproduce the single-GPU solver bit for bit; the multi-rank test runs
tests/dist_worker.py under torchrun when the box has >= 2 GPUs (peer stores over NVLink, epoch flags)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _system(grid=14):
    from learningsparsepreconditioner4gpu_b200 import systems
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    s = systems.poisson_sample(grid, seed=2, device="cuda")
    ws = systems.default_workspace(s)
    n = s.num_nodes
    A = to_csr_device(s.edge_index, s.matrix_values, n, s.mask)
    G = ws.inference_step_device(s)
    rhs = A.matvec(s.mask.reshape(-1).double())
    return s, ws, A, G, rhs


@pytest.mark.parametrize("mode", [1, 2, 4])
@pytest.mark.parametrize("precond", ["ext_spai", "ext_spai_scaled", "none", "diagonal"])
def test_world1_reproduces_single_gpu_solver(precond, mode):
    from learningsparsepreconditioner4gpu_b200 import parallel as par
    from learningsparsepreconditioner4gpu_b200.device import DevicePCG

    if mode == 4 and not precond.startswith("ext_spai"):
        pytest.skip("the fused persistent kernel is the SPAI iteration")
    s, ws, A, G, rhs = _system()
    n = s.num_nodes
    starts = par.partition_rows(par.row_weights_from_edges(s.edge_index, n), 1)
    assert starts == [0, n]
    spai = precond.startswith("ext_spai")
    d = par.DistPCG(A, G if spai else None, ws.epsilon, starts, par.extended_range(s.edge_index, 0, n), rank=0, world=1, mode=mode)
    x, res, hist = d.solve(rhs, precond, rtol=1e-8, history=True)
    ref = DevicePCG(A, mode=mode)
    if spai:
        ref.set_spai(G, ws.epsilon)
    x1, res1, hist1 = ref.solve(rhs, precond, rtol=1e-8, history=True)
    assert res.converged == 1 and res.iterations == res1.iterations
    assert torch.equal(hist, hist1)
    assert torch.equal(x, x1)
    assert d.info() == {"n_push": 0, "n_recv": 0, "push_elems": 0, "owned_offset": 0}
    d.close()


def test_slice_rows_matches_scipy():
    import ctypes as C

    from learningsparsepreconditioner4gpu_b200 import _lib
    from learningsparsepreconditioner4gpu_b200.device import DeviceCSR, _stream

    s, ws, A, G, rhs = _system(10)
    As = A.to_scipy()
    lib = _lib.load()
    h = C.c_void_p()
    r0, r1, shift, ncl = 200, 611, 57, 900
    _lib.check(lib.b200spai_csr_slice_rows(A._h, r0, r1, shift, ncl, _stream(), C.byref(h)), "slice")
    L = DeviceCSR(h).to_scipy()
    ref = As[r0:r1, shift:shift + ncl]
    assert L.shape == ref.shape and (abs(L - ref)).nnz == 0 and L.nnz == As[r0:r1].nnz
    # a window that does not contain every column is rejected
    rc = lib.b200spai_csr_slice_rows(A._h, r0, r1, 300, 100, _stream(), C.byref(h))
    assert rc != 0


def test_zero_rhs_and_repeat_solves():
    from learningsparsepreconditioner4gpu_b200 import parallel as par

    s, ws, A, G, rhs = _system(10)
    n = s.num_nodes
    d = par.DistPCG(A, G, ws.epsilon, [0, n], par.extended_range(s.edge_index, 0, n), rank=0, world=1)
    x, res, _ = d.solve(torch.zeros_like(rhs), "ext_spai")
    assert res.iterations == 0 and res.converged == 1 and float(x.abs().max()) == 0.0
    x1, r1, _ = d.solve(rhs, "ext_spai")
    x2, r2, _ = d.solve(rhs, "ext_spai")
    assert r1.iterations == r2.iterations and torch.equal(x1, x2)
    d.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("mode", [1, 2, 4])
def test_multi_rank_matches_single_gpu(mode):
    world = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29611 + mode), os.path.join(ROOT, "tests", "dist_worker.py"), "20", str(mode)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["world"] == world and out["ext_spai"]["n_push"] >= 1


@pytest.mark.parametrize("kind", ["poisson", "elast"])
def test_partitioned_system_world1_matches_global(kind):
    """PartitionedSystem with one rank: stage-wise GNN, local assembly (incl. G^T from the swapped edge list) and the
    solver built from pre-assembled local matrices reproduce the global path bit for bit."""
    from learningsparsepreconditioner4gpu_b200 import parallel as par, systems
    from learningsparsepreconditioner4gpu_b200.device import DevicePCG
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    if kind == "elast":
        s, b = systems.elasticity_sample(8, 4, 4, seed=1, device="cuda"), 3
    else:
        s, b = systems.poisson_sample(12, seed=2, device="cuda"), 1
    ws = systems.default_workspace(s, block_size=b)
    N = s.num_nodes
    n = N * b
    ps = par.PartitionedSystem(s.edge_index, N, b, [0, N], rank=0, world=1)
    g_t = ps.gnn_forward(ws.gnn, s.x, s.edge_attr)
    g_ref = ws(s.x, s.edge_index, s.edge_attr)
    assert torch.equal(g_t.reshape(-1, b, b), g_ref)
    A_loc, G_loc, GT_loc = ps.assemble(s.matrix_values, g_t, s.mask)
    A = to_csr_device(s.edge_index, s.matrix_values, n, s.mask)
    G = to_csr_device(s.edge_index, g_ref, n, s.mask)
    for loc, glob in ((A_loc, A.to_scipy()), (G_loc, G.to_scipy()), (GT_loc, G.to_scipy().T.tocsr())):
        glob.sort_indices()
        l = loc.to_scipy()
        assert np.array_equal(l.indptr, glob.indptr) and np.array_equal(l.indices, glob.indices) and np.array_equal(l.data, glob.data)
    rhs = A.matvec(s.mask.reshape(-1).double())
    d = ps.make_solver((A_loc, G_loc, GT_loc), ws.epsilon)
    x, res, hist = d.solve(rhs, "ext_spai", rtol=1e-8, history=True)
    ref = DevicePCG(A, mode=2)          # the 5-kernel driver the distributed loop is built from (auto would pick mode 3 here)
    ref.set_spai(G, ws.epsilon)
    x1, res1, hist1 = ref.solve(rhs, "ext_spai", rtol=1e-8, history=True)
    assert res.iterations == res1.iterations and torch.equal(hist, hist1) and torch.equal(x, x1)
    d.close()
    ps.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("kind", ["poisson", "elast"])
def test_multi_rank_partitioned_gnn_and_assembly(kind):
    world = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29631 + (kind == "elast")), os.path.join(ROOT, "tests", "dist_gnn_worker.py"), kind,
           "20" if kind == "poisson" else "24"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert out["world"] == world and out["sent"] > 0 and out["edges_gnn"] < out["E"]


def test_connection_reuse_across_systems():
    """close() with reuse parks the connected handle; the next solver of another size relayouts it and gives the same
    bits as a fresh one."""
    from learningsparsepreconditioner4gpu_b200 import parallel as par

    out = []
    for grid in (12, 9, 13):       # 13^3 exceeds the parked capacity (12^3 * 1.25): handle is rebuilt
        s, ws, A, G, rhs = _system(grid)
        n = s.num_nodes
        ext = par.extended_range(s.edge_index, 0, n)
        d = par.DistPCG(A, G, ws.epsilon, [0, n], ext, rank=0, world=1, reuse=True)
        x, res, hist = d.solve(rhs, "ext_spai", rtol=1e-8, history=True)
        d.close()
        f = par.DistPCG(A, G, ws.epsilon, [0, n], ext, rank=0, world=1)
        x1, res1, hist1 = f.solve(rhs, "ext_spai", rtol=1e-8, history=True)
        f.close()
        assert res.iterations == res1.iterations and torch.equal(hist, hist1) and torch.equal(x, x1)
        out.append(len(par.DistPCG._connections))
    assert out == [1, 1, 1]
    par.DistPCG.drop_connections()
    assert not par.DistPCG._connections
