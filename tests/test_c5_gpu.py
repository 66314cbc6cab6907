"""GPU parity tests of the code path the bench TIMES (BASELINE config C5 family), through the C ABI.

The small goldens (<= 400 rows) never reach it: above 2^18 stored entries the SpMV is the TMA-staged streaming kernel
(lossless fp32-value / 16-bit-column-step streams, fused AXPY_DOT / DOT / SCALE epilogues), above 131 072 rows the solve
is driven by the CUDA-graph WHILE loop (mode 2) or by the fused persistent kernel (mode 4).  Here those run against

  * tests/golden/c5_64.npz / c5_128.npz -- minted by tests/golden/make_golden_c5.py through the UNMODIFIED reference
    modules (workspace.forward, to_csr_cpu, get_pcg_iter_time_scipy): iteration count, residual history, samples and
    checksums of G and of the solution for the jittered Kuhn-tet Poisson systems bench.py solves (seed 100);
  * the oracle (oracle/gnn.py, oracle/pcg.py) recomputed on the box's host cores, itself checked against the fixture's
    G samples (that pins the oracle to the reference at size);
  * scipy directly, for the operator z = G (G^T r) [/ d] + eps r [/ d] and for the fallback streams (fp64 values, int32
    columns).

Bars (stated where asserted): G within 1e-5 of max|G|; history entries 0..5 to 1e-12 on identical G; solution samples to
1e-4 of max|x|; iteration count on the REFERENCE's G (oracle-recomputed) and on the CUDA G within IT_TOL of the
reference's: +-1 at 64^3 (234 iterations); at 128^3 (522 iterations) the reference's own count is only defined up to
-3 .. +7 -- that is how far its recurrence moves when nothing but the summation order of its dot products changes
(scripts/cg_order_sensitivity.py -> profiles/r02_cg_order_sensitivity.jsonl: 529 / 519 / 520 against 522, histories apart
by > 1 % from iteration 23) -- so the stated tolerance there is +-7 and the measured delta is printed (-2 .. 0)."""
import os

import numpy as np
import pytest
import scipy.sparse as sp
import torch

from conftest import GOLDEN_DIR
from test_pcg_gpu import history_within_envelope

pytestmark = pytest.mark.gpu

MODES = (2, 4)  # graph WHILE driver, fused persistent kernel
IT_TOL = {64: 1, 128: 7}


def _close(a, b, rel):
    return abs(float(a) - float(b)) <= rel * max(abs(float(a)), abs(float(b)), 1e-300)


def _load(grid):
    return np.load(os.path.join(GOLDEN_DIR, f"c5_{grid}.npz"), allow_pickle=False)


def _system(grid, g):
    """The fixture's system regenerated on the CPU exactly as make_golden_c5.py did, its inputs proven equal by checksum."""
    from learningsparsepreconditioner4gpu_b200 import systems

    sample = systems.poisson_sample(grid, seed=int(g["seed"]), gen_device="cpu", device="cpu")
    assert sample.num_nodes == int(g["N"]) and sample.num_edges == int(g["E"])
    for name, t in (("in_x", sample.x), ("in_edge_attr", sample.edge_attr), ("in_mask", sample.mask)):
        a = t.numpy().astype(np.float64).reshape(-1)
        assert _close(np.abs(a).sum(), g[name + "_abs_sum"], 1e-9), name
        assert _close(np.sqrt((a * a).sum()), g[name + "_l2"], 1e-9), name
    ws = systems.default_workspace(sample)
    sd = {str(k): torch.from_numpy(g["sd/" + str(k)].copy()) for k in g["sd_keys"]}
    ws.load_state_dict(sd)                       # the fixture's weights (equal to the seed-42 init; pinned elsewhere)
    return sample, ws, sd


def _solve_and_check(g, A, G, b, eps, mode, same_G):
    from learningsparsepreconditioner4gpu_b200 import device as d

    solver = d.DevicePCG(A, mode=mode)
    solver.set_spai(G, eps)
    x, res, hist = solver.solve(b, "ext_spai", rtol=1e-6, history=True)
    assert res.reserved == mode, (res.reserved, mode)            # the driver that really ran
    assert res.converged == 1 and res.breakdown == 0
    ref_it, ref_hist = int(g["iters_pcg"]), g["res_hist"]
    dit = res.iterations - ref_it
    print(f"[c5 grid={int(g['grid'])} mode={mode} {'reference G' if same_G else 'CUDA G'}] iterations {res.iterations} "
          f"vs reference {ref_it} (delta {dit:+d})")
    assert abs(dit) <= IT_TOL[int(g["grid"])], (res.iterations, ref_it)
    h = hist.cpu().numpy()
    assert np.allclose(h[:6], ref_hist[:6], rtol=1e-12 if same_G else 1e-4)
    if ref_it <= 250:      # beyond that two correctly rounded runs are > 10 % apart entry by entry (see the module docstring)
        assert history_within_envelope(h, ref_hist, shift=2, slack=0.25) == []
    assert h[-1] < 1e-6 * h[0] <= h[-2]
    xs = x.cpu().numpy()
    assert np.allclose(xs[g["x_sample_ids"]], g["x_samples"], rtol=0, atol=1e-4 * float(g["xsol_max_abs"]))
    r = b - A.matvec(x)
    assert float(r.norm() / b.norm()) < 2e-6
    return res, xs


@pytest.mark.parametrize("grid", [64, 128])
def test_c5_golden_gnn_and_pcg_on_the_timed_path(grid):
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    g = _load(grid)
    sample, ws, sd = _system(grid, g)
    eps = float(g["epsilon"])
    n = sample.num_nodes
    dev = sample.to("cuda")
    ws = ws.cuda()
    # ---- GNN: CUDA G against the reference's G (10 000 seeded samples + checksums) ----
    with torch.no_grad():
        Gv = ws(dev.x, dev.edge_index, dev.edge_attr).reshape(-1)
    gmax = float(g["g_max_abs"])
    got = Gv[torch.from_numpy(g["g_sample_ids"]).cuda()].cpu().numpy()
    dG = float(np.abs(got - g["g_samples"]).max())
    print(f"[c5 grid={grid}] max |dG| / max|G| on the samples = {dG / gmax:.3e}")
    assert dG <= 1e-5 * gmax
    assert _close(Gv.double().abs().sum().item(), g["g_abs_sum"], 1e-5)
    assert _close(Gv.double().norm().item(), g["g_l2"], 1e-5)
    # ---- assembly: same pattern sizes as the reference's to_csr_cpu ----
    A = to_csr_device(dev.edge_index, dev.matrix_values, n, dev.mask)
    G = to_csr_device(dev.edge_index, Gv.reshape(-1, 1, 1), n, dev.mask)
    assert A.nnz == int(g["nnzA"]) and G.nnz == int(g["nnzG"])
    assert A.nnz >= (1 << 18) and n > 131072
    assert A.stream_bytes() == (4, 2) and G.stream_bytes() == (4, 2)     # fp32 values + 16-bit column steps
    b = A.matvec(dev.mask.reshape(-1).double())
    assert _close(b.norm().item(), g["b_norm"], 1e-12)
    # ---- PCG on the CUDA G, both large-system drivers ----
    for mode in MODES:
        _solve_and_check(g, A, G, b, eps, mode, same_G=False)
    # ---- PCG on the REFERENCE's G: the oracle forward recomputed on the host cores, pinned by the fixture's samples ----
    from oracle import gnn as ognn

    Go = ognn.workspace_forward(sd, sample.x, sample.edge_index, sample.edge_attr, 1).reshape(-1)
    assert np.abs(Go.numpy()[g["g_sample_ids"]] - g["g_samples"]).max() <= 1e-7    # chunk/thread-boundary erf noise only
    Gref = to_csr_device(dev.edge_index, Go.cuda().reshape(-1, 1, 1), n, dev.mask)
    for mode in MODES:
        _solve_and_check(g, A, Gref, b, eps, mode, same_G=True)


def _spd_pair(n, per_row, seed, fp32_values, narrow):
    """(A SPD, G) with nnz >= 2^18: `narrow` keeps columns within +-3000 of the diagonal (16-bit steps), otherwise they
    are spread over the whole row (steps > 65535 -> int32 column stream); fp32_values=False -> fp64 value stream."""
    rng = np.random.default_rng(seed)
    rows = np.repeat(np.arange(n), per_row)
    if narrow:
        cols = np.clip(rows + rng.integers(-3000, 3001, rows.size), 0, n - 1)
    else:
        cols = rng.integers(0, n, rows.size)
    vals = rng.standard_normal(rows.size) * 0.1
    M = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()
    M = M + M.T
    d = np.asarray(abs(M).sum(axis=1)).reshape(-1) + 1.0
    A = (M + sp.diags(d)).tocsr()
    G = (sp.tril(A, format="csr") @ sp.diags(1.0 / np.sqrt(d))).tocsr()
    G.data *= 1.0 + 0.1 * rng.standard_normal(G.nnz)
    if fp32_values:
        A.data = A.data.astype(np.float32).astype(np.float64)
        G.data = G.data.astype(np.float32).astype(np.float64)
    A.sort_indices()
    G.sort_indices()
    return A, G


@pytest.mark.parametrize("fp32_values,narrow", [(True, True), (False, False), (True, False), (False, True)])
def test_stream_epilogues_and_fallback_streams_vs_scipy(fp32_values, narrow):
    """The streaming kernel's fused epilogues (AXPY_DOT with eps, DOT, SCALE) at nnz >= 2^18 against scipy, in all four
    stream formats, and a whole PCG solve in the same formats against the oracle's scipy-cg recurrence."""
    from learningsparsepreconditioner4gpu_b200 import device as d
    from oracle import pcg as opcg

    n = 150000
    A, G = _spd_pair(n, 4, 7, fp32_values, narrow)
    assert A.nnz >= (1 << 18) and G.nnz >= (1 << 18)
    Ad, Gd = d.DeviceCSR.from_scipy(A), d.DeviceCSR.from_scipy(G)
    want = (4 if fp32_values else 8, 2 if narrow else 4)
    assert Ad.stream_bytes() == want and Gd.stream_bytes() == want
    eps = 3e-3
    rng = np.random.default_rng(1)
    r = rng.standard_normal(n)
    GT = sp.csr_matrix(G.T)
    GT.sort_indices()
    dd = A.diagonal()
    solver = d.DevicePCG(Ad, mode=2)
    solver.set_spai(Gd, eps)
    z = solver.apply_precond(torch.from_numpy(r).cuda(), "ext_spai").cpu().numpy()
    assert np.array_equal(z, G @ (GT @ r) + eps * r)                      # same roundings as scipy: bit-exact
    zs = solver.apply_precond(torch.from_numpy(r).cuda(), "ext_spai_scaled").cpu().numpy()
    assert np.allclose(zs, G @ ((GT @ r) / dd) + eps * r / dd, rtol=1e-14, atol=0)
    b = A @ rng.standard_normal(n)
    bd = torch.from_numpy(b).cuda()
    for kind in ("ext_spai", "ext_spai_scaled"):
        xo, its, ohist = opcg.pcg_restated(A, b, opcg.make_preconditioner(kind, A, G, eps), rtol=1e-8)
        for mode in MODES:
            s2 = d.DevicePCG(Ad, mode=mode)
            s2.set_spai(Gd, eps)
            x, res, hist = s2.solve(bd, kind, rtol=1e-8, history=True)
            assert res.reserved == mode and res.converged == 1
            h = hist.cpu().numpy()
            # entries 0..3 pin the arithmetic of the epilogues in every format
            assert np.allclose(h[:4], ohist[:4], rtol=1e-12), (kind, mode, h[:6], ohist[:6])
            if narrow:
                # banded pair: non-monotone history with sharp dips; the REFERENCE recurrence itself moves by up to +3
                # iterations and by factors of 2-10 in the history when only its dot-product order changes
                # (scripts/band_pair_sensitivity.py, CPU), identically for the stream, graph and fused drivers
                # (scripts/dbg_epilogue_hist.py) -- count +-3, no envelope
                assert abs(res.iterations - its) <= 3, (kind, mode, res.iterations, its)
            else:
                assert abs(res.iterations - its) <= 1, (kind, mode, res.iterations, its)
                assert np.allclose(h[:6], ohist[:6], rtol=1e-12)
                assert history_within_envelope(h, ohist, shift=2, slack=0.25) == []
            assert np.allclose(x.cpu().numpy(), xo, rtol=0, atol=1e-6 * np.abs(xo).max())


def test_fused_kernel_edge_cases():
    """mode 4: zero rhs, max_iter cut-off, determinism, profile hook; auto mode keeps the graph driver on one GPU."""
    from learningsparsepreconditioner4gpu_b200 import device as d

    A, G = _spd_pair(140000, 4, 3, True, True)
    Ad, Gd = d.DeviceCSR.from_scipy(A), d.DeviceCSR.from_scipy(G)
    b = torch.from_numpy(A @ np.random.default_rng(2).standard_normal(A.shape[0])).cuda()
    solver = d.DevicePCG(Ad, mode=4)
    solver.set_spai(Gd, 3e-3)
    x, res, hist = solver.solve(torch.zeros_like(b), "ext_spai", history=True)
    assert res.iterations == 0 and res.converged == 1 and float(x.abs().max()) == 0.0 and hist.shape[0] == 1
    x, res, hist = solver.solve(b, "ext_spai", rtol=1e-14, max_iter=5, history=True)
    assert res.iterations == 5 and res.converged == 0 and hist.shape[0] == 6
    x1, r1, h1 = solver.solve(b, "ext_spai", rtol=1e-8, history=True)
    x2, r2, h2 = solver.solve(b, "ext_spai", rtol=1e-8, history=True)
    assert r1.iterations == r2.iterations and torch.equal(x1, x2) and torch.equal(h1, h2)
    ms, its = solver.fused_profile()
    assert its == r2.iterations and all(m > 0 for m in ms)
    ref = d.DevicePCG(Ad, mode=2)
    ref.set_spai(Gd, 3e-3)
    x3, r3, h3 = ref.solve(b, "ext_spai", rtol=1e-8, history=True)
    assert abs(r3.iterations - r1.iterations) <= 3 and torch.allclose(h3[:4], h1[:4], rtol=1e-12, atol=0)   # banded pair: see above
    auto = d.DevicePCG(Ad)
    auto.set_spai(Gd, 3e-3)
    _, ra, _ = auto.solve(b, "ext_spai", rtol=1e-8)
    assert ra.reserved == 2 and ra.converged == 1
    with pytest.raises(Exception):
        solver.solve(b, "diagonal", rtol=1e-8)                      # mode 4 is the SPAI iteration
