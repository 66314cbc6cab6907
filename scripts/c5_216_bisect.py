"""Which piece of the path departs from the oracle at 216^3?  (measurement / debugging, GPU box)

    python scripts/c5_216_bisect.py [grid]

Builds the fixture system on the CPU, a value-independent stand-in for G on A's pattern, and compares piece by piece:
device assembly vs oracle.assemble (pattern + values), device SpMV / transpose / preconditioner apply vs scipy on random
vectors, and the first PCG iterations (device history vs oracle.pcg.pcg_restated)."""
import json
import os
import sys
import time

import numpy as np
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    grid = int(sys.argv[1]) if len(sys.argv) > 1 else 216
    real_g = len(sys.argv) > 2 and sys.argv[2] == "--real-g"
    from learningsparsepreconditioner4gpu_b200 import device as d, systems
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device
    from oracle import assemble as oasm, pcg as opcg

    out = {"grid": grid}
    t0 = time.time()
    s = systems.poisson_sample(grid, seed=100, gen_device="cpu", device="cpu")
    n, E = s.num_nodes, s.num_edges
    rng = np.random.default_rng(0)
    gvals = torch.from_numpy((0.1 + 0.5 * rng.standard_normal(E)).astype(np.float32))
    out["gen_s"] = round(time.time() - t0, 1)
    fx = None
    if real_g:   # the oracle's G of the fixture weights, and every checksum the fixture holds
        from oracle import gnn as ognn

        fx = np.load(os.path.join(ROOT, "tests", "golden", f"c5_{grid}.npz"), allow_pickle=False)
        sd = {str(k): torch.from_numpy(fx["sd/" + str(k)].copy()) for k in fx["sd_keys"]}
        t0 = time.time()
        gvals = ognn.workspace_forward_chunked(sd, s.x, s.edge_index, s.edge_attr, 1).reshape(-1)
        out["oracle_g_s"] = round(time.time() - t0, 1)

        def cs(a):
            a = np.asarray(a, dtype=np.float64).reshape(-1)
            return {"sum": float(a.sum()), "abs_sum": float(np.abs(a).sum()), "l2": float(np.sqrt((a * a).sum())), "max_abs": float(np.abs(a).max())}

        for name, arr in (("g", gvals.numpy()), ("in_x", s.x.numpy()), ("in_edge_attr", s.edge_attr.numpy()), ("in_mask", s.mask.numpy())):
            c = cs(arr)
            out["checksum_rel_err_" + name] = {k: abs(c[k] - float(fx[f"{name}_{k}"])) / max(abs(float(fx[f"{name}_{k}"])), 1e-300) for k in c}
        out["g_samples_max_abs_diff"] = float(np.abs(gvals.numpy()[fx["g_sample_ids"]] - fx["g_samples"]).max())
    t0 = time.time()
    A = oasm.to_csr_cpu_sorted(s.edge_index, s.matrix_values, n, s.mask)
    G = oasm.to_csr_cpu_sorted(s.edge_index, gvals, n, s.mask)
    out["oracle_assemble_s"] = round(time.time() - t0, 1)
    dev = s.to("cuda")
    Ad = to_csr_device(dev.edge_index, dev.matrix_values, n, dev.mask)
    Gd = to_csr_device(dev.edge_index, gvals.cuda().reshape(-1, 1, 1), n, dev.mask)
    for name, M, Md in (("A", A, Ad), ("G", G, Gd)):
        L = Md.to_scipy()
        out[name + "_pattern_equal"] = bool(np.array_equal(L.indptr, M.indptr) and np.array_equal(L.indices, M.indices))
        out[name + "_values_equal"] = bool(np.array_equal(L.data, M.data))
        out[name + "_nnz"] = [int(L.nnz), int(M.nnz)]
        del L
    x = rng.standard_normal(n)
    xd = torch.from_numpy(x).cuda()
    for name, M, Md in (("A", A, Ad), ("G", G, Gd)):
        y = Md.matvec(xd).cpu().numpy()
        ref = M @ x
        out[name + "_spmv_max_abs_diff"] = float(np.abs(y - ref).max())
        out[name + "_spmv_equal"] = bool(np.array_equal(y, ref))
    GT = sp.csr_matrix(G.T)
    GT.sort_indices()
    GTd = Gd.transpose()
    yt = GTd.matvec(xd).cpu().numpy()
    out["GT_spmv_equal"] = bool(np.array_equal(yt, GT @ x))
    out["GT_spmv_max_abs_diff"] = float(np.abs(yt - GT @ x).max())
    eps = 3e-3
    solver = d.DevicePCG(Ad, mode=2)
    solver.set_spai(Gd, eps)
    z = solver.apply_precond(xd, "ext_spai").cpu().numpy()
    zref = G @ (GT @ x) + eps * x
    out["precond_equal"] = bool(np.array_equal(z, zref))
    out["precond_max_rel_diff"] = float(np.abs(z - zref).max() / np.abs(zref).max())
    b = A @ s.mask.numpy().reshape(-1).astype(np.float64)
    bd = Ad.matvec(dev.mask.reshape(-1).double())
    out["b_equal"] = bool(np.array_equal(bd.cpu().numpy(), b))
    K = 8
    t0 = time.time()
    _, _, oh = opcg.pcg_restated(A, b, opcg.make_preconditioner("ext_spai", A, G, eps), rtol=1e-30, max_iter=K)
    out["oracle_pcg_s"] = round(time.time() - t0, 1)
    if fx is not None:
        c = {"sum": float(A.data.sum()), "abs_sum": float(np.abs(A.data).sum()), "l2": float(np.sqrt((A.data * A.data).sum())), "max_abs": float(np.abs(A.data).max())}
        out["checksum_rel_err_a_data"] = {k: abs(c[k] - float(fx[f"a_data_{k}"])) / abs(float(fx[f"a_data_{k}"])) for k in c}
        out["box_oracle_vs_fixture_hist"] = [float(abs(a - b_) / b_) for a, b_ in zip(oh, fx["res_hist"])]
        out["fixture_nnz"] = [int(fx["nnzA"]), int(fx["nnzG"])]
    for mode in (2, 4):
        sv = d.DevicePCG(Ad, mode=mode)
        sv.set_spai(Gd, eps)
        _, res, h = sv.solve(bd, "ext_spai", rtol=1e-30, max_iter=K, history=True)
        h = h.cpu().numpy()
        out[f"hist_rel_dev_mode{mode}"] = [float(abs(a - b_) / b_) for a, b_ in zip(h, oh)]
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
