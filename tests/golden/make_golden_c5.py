"""Mint the LARGE-system fixtures tests/golden/c5_<grid>.npz (BASELINE config C5 family: jittered Kuhn-tet Poisson).

Run once in the build container (where /root/reference exists):

    python tests/golden/make_golden_c5.py 64 128          # through the UNMODIFIED reference modules (ref_shim)
    python tests/golden/make_golden_c5.py --oracle 216    # the oracle (chunked GNN + restated scipy cg) at bench size

Why: the small goldens (<= 400 rows) never reach the kernels the bench times (streaming SpMV with fused epilogues,
graph-WHILE driver, nnz >= 2^18 and > 131072 rows).  These fixtures pin that path at 64^3 (262 k rows, 3.9 M stored
entries), 128^3 (2.1 M rows) and at the bench size 216^3 (10.08 M rows).

A full G at these sizes is 15 MB .. 600 MB, so the fixture stores instead
  * iteration count of the reference's get_pcg_iter_time_scipy (neural_cg/utils/validate.py:163-201),
  * the recurrence-residual history ||r_k|| of the same solve (oracle.pcg.pcg_restated: the scipy cg recurrence written out;
    its count is asserted equal to the reference function's),
  * checksums and 10 000 seeded samples of G (reference workspace forward, neural_cg/workspace.py:92-94),
  * checksums and samples of the solution x,
  * checksums of the INPUTS (x, edge_attr, mask), so that a test regenerating the system with
    systems.poisson_sample(grid, seed=100, gen_device="cpu") can prove it solves the same system.
The tests recompute G with the oracle on the box's CPU, check it against the fixture's samples/checksums (that pins the
oracle to the reference at size), and compare the CUDA path with both.

At 216^3 the reference modules would need 2 x 29 GB per [E,48] concatenation, so `--oracle` runs the oracle's
edge-chunked forward (oracle.gnn.workspace_forward_chunked; this script measures its distance to the reference forward at 64^3 and
128^3: a handful of entries move by <= 1e-7 because torch's CPU GELU uses the scalar erf at thread-range ends) and pcg_restated only.  The fixture records which producer made it.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

SEED = 100          # bench.py's system seed on rank 0
N_SAMPLES = 10000
EPS = 3.0e-3


def checksums(a: np.ndarray, prefix: str):
    a64 = np.asarray(a, dtype=np.float64).reshape(-1)
    return {prefix + "_sum": np.float64(a64.sum()), prefix + "_abs_sum": np.float64(np.abs(a64).sum()),
            prefix + "_l2": np.float64(np.sqrt((a64 * a64).sum())), prefix + "_max_abs": np.float64(np.abs(a64).max())}


def sample_ids(count: int, total: int) -> np.ndarray:
    rng = np.random.default_rng(12345)
    return np.sort(rng.choice(total, size=min(count, total), replace=False)).astype(np.int64)


def mint(grid: int, use_reference: bool):
    from learningsparsepreconditioner4gpu_b200 import systems
    from oracle import assemble as oasm, gnn as ognn, pcg as opcg

    t0 = time.time()
    sample = systems.poisson_sample(grid, seed=SEED, gen_device="cpu", device="cpu")
    N, E = sample.num_nodes, sample.num_edges
    print(f"[{grid}] generated N={N} E={E} in {time.time() - t0:.1f}s", flush=True)
    ws_sd = None
    t0 = time.time()
    if use_reference:
        import ref_shim

        ref_shim.install("/root/reference")
        import make_golden as mg  # noqa: E402  builds the reference workspace with the resolved default config
        from neural_cg.utils import validate as rv  # reference

        ws = mg.build_workspace(sample.x.shape[1], sample.edge_attr.shape[1], 1, seed=42)
        ws_sd = {k: v.detach().clone() for k, v in ws.state_dict().items()}
        with torch.no_grad():
            G = ws(sample.x, sample.edge_index, sample.edge_attr)
        producer = "reference modules via ref_shim (workspace.forward, to_csr_cpu, get_pcg_iter_time_scipy)"
        # the oracle, un-chunked and chunked, must reproduce it bit for bit
        G_o = ognn.workspace_forward(ws_sd, sample.x, sample.edge_index, sample.edge_attr, 1)
        assert torch.equal(G_o, G), "oracle forward differs from the reference at size"
        # the chunked forward runs the same kernels on the same rows; torch's CPU GELU evaluates the few elements at the
        # end of each thread's range with the scalar erf instead of the vector one, and where those ranges end depends on
        # the tensor length -- so chunking moves a handful of entries by one fp32 ulp of an intermediate (measured here)
        G_c = ognn.workspace_forward_chunked(ws_sd, sample.x, sample.edge_index, sample.edge_attr, 1, chunk=1 << 20)
        d_chunk = float((G_c - G).abs().max())
        print(f"[{grid}] chunked oracle vs reference: max |dG| = {d_chunk:.3e} ({int((G_c != G).sum())} entries differ)", flush=True)
        assert d_chunk <= 1e-7, "chunked oracle forward differs from the reference at size"
        del G_o, G_c
    else:
        ws = systems.default_workspace(sample)     # same seed-42 init as the reference (tests/test_host_logic.py pins it)
        ws_sd = {k: v.detach().clone() for k, v in ws.state_dict().items()}
        G = ognn.workspace_forward_chunked(ws_sd, sample.x, sample.edge_index, sample.edge_attr, 1, chunk=1 << 22)
        producer = "oracle (chunked forward, to_csr_cpu_sorted, pcg_restated); reference modules do not fit host memory at this size"
    print(f"[{grid}] G in {time.time() - t0:.1f}s  max|G|={float(G.abs().max()):.5f}", flush=True)
    n = N
    t0 = time.time()
    if use_reference:
        A = rv.to_csr_cpu(sample.edge_index, sample.matrix_values, n, sample.mask)
        Gc = rv.to_csr_cpu(sample.edge_index, G, n, sample.mask)
    else:
        A = oasm.to_csr_cpu_sorted(sample.edge_index, sample.matrix_values, n, sample.mask)
        Gc = oasm.to_csr_cpu_sorted(sample.edge_index, G.reshape(-1), n, sample.mask)
    gt = sample.mask.numpy().reshape(-1).astype(np.float64)
    print(f"[{grid}] assembled nnzA={A.nnz} nnzG={Gc.nnz} in {time.time() - t0:.1f}s", flush=True)
    t0 = time.time()
    it_ref = -1
    if use_reference:
        it_ref = int(rv.get_pcg_iter_time_scipy(A, gt, Gc, EPS, rtol=1e-6))
        print(f"[{grid}] reference get_pcg_iter_time_scipy: {it_ref} iterations in {time.time() - t0:.1f}s", flush=True)
        t0 = time.time()
    b = A @ gt
    xs, its, hist = opcg.pcg_restated(A, b, opcg.make_preconditioner("ext_spai", A, Gc, EPS), rtol=1e-6)
    print(f"[{grid}] pcg_restated: {its} iterations in {time.time() - t0:.1f}s", flush=True)
    if use_reference:
        assert its == it_ref, (its, it_ref)
    true_res = float(np.linalg.norm(b - A @ xs))
    ids_e = sample_ids(N_SAMPLES, E)
    ids_n = sample_ids(N_SAMPLES, n)
    Gn = G.numpy().reshape(-1)
    out = dict(
        grid=np.int64(grid), seed=np.int64(SEED), N=np.int64(N), E=np.int64(E), epsilon=np.float64(EPS),
        producer=np.array(producer), nnzA=np.int64(A.nnz), nnzG=np.int64(Gc.nnz),
        iters_pcg=np.int64(its), iters_reference_fn=np.int64(it_ref), res_hist=np.asarray(hist, dtype=np.float64),
        b_norm=np.float64(np.linalg.norm(b)), true_residual=np.float64(true_res),
        g_sample_ids=ids_e, g_samples=Gn[ids_e].astype(np.float32),
        x_sample_ids=ids_n, x_samples=xs[ids_n].astype(np.float64),
        sd_keys=np.array(list(ws_sd.keys())),
    )
    for k, v in ws_sd.items():
        out["sd/" + k] = v.numpy()
    out.update(checksums(Gn, "g"))
    out.update(checksums(xs, "xsol"))
    out.update(checksums(sample.x.numpy(), "in_x"))
    out.update(checksums(sample.edge_attr.numpy(), "in_edge_attr"))
    out.update(checksums(sample.mask.numpy(), "in_mask"))
    out.update(checksums(A.data, "a_data"))
    path = os.path.join(HERE, f"c5_{grid}.npz")
    np.savez_compressed(path, **out)
    print(f"[{grid}] iters={its} true_res/|b|={true_res / np.linalg.norm(b):.3e} -> {path} "
          f"({os.path.getsize(path) / 1024:.0f} KiB)", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("grids", type=int, nargs="+")
    ap.add_argument("--oracle", action="store_true", help="use the oracle instead of the reference modules (216^3)")
    a = ap.parse_args()
    for g in a.grids:
        mint(g, not a.oracle)


if __name__ == "__main__":
    main()
