"""Native 3x3 block stream (csrc/bsr.cuh) for the elasticity shape (BASELINE config C3, `data.block_size=3`): the reference
keeps b x b blocks (neural_cg/data.py:134-156, neural_cg/utils/validate.py:36-45); here a matrix assembled from [E,3,3]
blocks is streamed as 9 fp32 values + one 16-bit block-column step per block.  Bar: every result carries the BITS of the
scalar path -- y = A x, the transposed product, the fused epilogues and whole solves equal scipy's csr_matvec of the
masked / zero-pruned matrix exactly (the entries pruned from the scalar pattern are explicit +0 in the blocks)."""
import numpy as np
import pytest
import scipy.sparse as sp
import torch

pytestmark = pytest.mark.gpu


def _elast(nx, ny, nz):
    from learningsparsepreconditioner4gpu_b200 import systems
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    s = systems.elasticity_sample(nx, ny, nz, seed=2, device="cuda")
    ws = systems.default_workspace(s, block_size=3)
    n = s.num_nodes * 3
    A = to_csr_device(s.edge_index, s.matrix_values, n, s.mask)
    with torch.no_grad():
        G = to_csr_device(s.edge_index, ws(s.x, s.edge_index, s.edge_attr), n, s.mask)
    return s, ws, A, G, n


def test_block_stream_is_used_and_is_bit_identical_to_scipy():
    s, ws, A, G, n = _elast(40, 14, 14)
    assert A.nnz >= (1 << 18)
    bs, nblocks = A.bsr_info()
    assert bs == 3 and nblocks > 0 and 38 * nblocks < 6 * A.nnz          # fewer bytes than the scalar stream
    print(f"[bsr] n = {n}, nnz(A) = {A.nnz}, blocks = {nblocks}: {38 * nblocks / (6 * A.nnz):.3f} of the scalar stream's bytes")
    As, Gs = A.to_scipy(), G.to_scipy()
    rng = np.random.default_rng(0)
    x = rng.standard_normal(n)
    xd = torch.from_numpy(x).cuda()
    assert np.array_equal(A.matvec(xd, kernel=2).cpu().numpy(), As @ x)
    assert np.array_equal(G.matvec(xd, kernel=2).cpu().numpy(), Gs @ x)
    GT = G.transpose()
    assert GT.bsr_info()[0] == 3
    GTs = sp.csr_matrix(Gs.T)
    GTs.sort_indices()
    assert np.array_equal(GT.matvec(xd, kernel=2).cpu().numpy(), GTs @ x)
    # the vector kernel (scalar CSR) agrees to rounding: the two layouts hold the same matrix
    yv = A.matvec(xd, kernel=1).cpu().numpy()
    assert np.allclose(yv, As @ x, rtol=1e-13, atol=1e-13 * np.abs(As @ x).max())


def test_block_stream_epilogues_and_solve_match_the_oracle():
    from learningsparsepreconditioner4gpu_b200 import device as d
    from oracle import pcg as opcg

    s, ws, A, G, n = _elast(40, 14, 14)
    As, Gs = A.to_scipy(), G.to_scipy()
    GTs = sp.csr_matrix(Gs.T)
    GTs.sort_indices()
    eps = ws.epsilon
    r = np.random.default_rng(1).standard_normal(n)
    solver = d.DevicePCG(A, mode=2)
    solver.set_spai(G, eps)
    z = solver.apply_precond(torch.from_numpy(r).cuda(), "ext_spai").cpu().numpy()
    assert np.array_equal(z, Gs @ (GTs @ r) + eps * r)                   # AXPY_DOT epilogue on blocks: scipy's bits
    dd = As.diagonal()
    zs = solver.apply_precond(torch.from_numpy(r).cuda(), "ext_spai_scaled").cpu().numpy()
    assert np.allclose(zs, Gs @ ((GTs @ r) / dd) + eps * r / dd, rtol=1e-14, atol=0)
    b = As @ s.mask.cpu().numpy().reshape(-1).astype(np.float64)
    bd = A.matvec(s.mask.reshape(-1).double())
    assert np.array_equal(bd.cpu().numpy(), b)
    xo, its, oh = opcg.pcg_restated(As, b, opcg.make_preconditioner("ext_spai", As, Gs, eps), rtol=1e-6)
    x, res, hist = solver.solve(bd, "ext_spai", rtol=1e-6, history=True)
    print(f"[bsr solve] iterations {res.iterations} vs oracle {its} (delta {res.iterations - its:+d})")
    assert res.reserved == 2 and res.converged == 1 and abs(res.iterations - its) <= 1
    assert np.allclose(hist.cpu().numpy()[:6], oh[:6], rtol=1e-12)
    assert np.abs(x.cpu().numpy() - xo).max() <= 1e-4 * np.abs(xo).max()


def test_scalar_matrices_and_sparse_blocks_keep_the_scalar_stream():
    from learningsparsepreconditioner4gpu_b200 import systems
    from learningsparsepreconditioner4gpu_b200.device import DeviceCSR
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    p = systems.poisson_sample(40, seed=1, device="cuda")
    A = to_csr_device(p.edge_index, p.matrix_values, p.num_nodes, p.mask)
    assert A.bsr_info() == (1, 0)
    # 3x3 blocks that are diagonal only: 3 of 9 entries -> 38 B per block > 18 B scalar, block stream not worth it
    e = systems.elasticity_sample(30, 10, 10, seed=1, device="cuda")
    vals = torch.zeros_like(e.matrix_values)
    for k in range(3):
        vals[:, k, k] = e.matrix_values[:, k, k].abs() + 1.0
    ones = torch.ones(e.num_nodes, 3, device="cuda")                 # a mask makes the assembly prune the exact zeros
    D = DeviceCSR.from_edges(e.edge_index, vals, e.num_nodes, ones)
    assert D.nnz == 3 * e.num_edges and D.bsr_info() == (1, 0)
    x = torch.randn(D.shape[1], dtype=torch.float64, device="cuda")
    assert np.array_equal(D.matvec(x, kernel=2).cpu().numpy(), D.to_scipy() @ x.cpu().numpy())
