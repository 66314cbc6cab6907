"""Measured iteration deltas of the CUDA path against the C5 fixtures (tests/golden/c5_*.npz), one JSON line per case.

    python scripts/c5_parity_probe.py 64 128 [--oracle-g] [--modes 2,4]

For every grid: CUDA G (with the GELU variant selected by B200SPAI_GNN_CFG in the environment) -> |dG| against the
fixture's samples, PCG iterations per driver; with --oracle-g also the oracle forward on the host cores (the reference's
G up to 1e-7, pinned by the same samples) -> PCG iterations on THAT G.  Measurement only (feeds DESIGN.md §4)."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("grids", type=int, nargs="+")
    ap.add_argument("--oracle-g", action="store_true")
    ap.add_argument("--modes", default="2,4")
    a = ap.parse_args()
    from learningsparsepreconditioner4gpu_b200 import device as d, systems
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    modes = [int(m) for m in a.modes.split(",")]
    for grid in a.grids:
        g = np.load(os.path.join(ROOT, "tests", "golden", f"c5_{grid}.npz"), allow_pickle=False)
        t0 = time.time()
        sample = systems.poisson_sample(grid, seed=int(g["seed"]), gen_device="cpu", device="cpu")   # as the fixture was minted
        ws = systems.default_workspace(sample)
        sd = {str(k): torch.from_numpy(g["sd/" + str(k)].copy()) for k in g["sd_keys"]}
        ws.load_state_dict(sd)
        ws = ws.cuda()
        dev = sample.to("cuda")
        n = sample.num_nodes
        with torch.no_grad():
            Gv = ws(dev.x, dev.edge_index, dev.edge_attr).reshape(-1)
        got = Gv[torch.from_numpy(g["g_sample_ids"]).cuda()].cpu().numpy()
        out = {"grid": grid, "gnn_cfg": os.environ.get("B200SPAI_GNN_CFG", "default"), "ref_iters": int(g["iters_pcg"]),
               "dG_over_maxG": float(np.abs(got - g["g_samples"]).max() / float(g["g_max_abs"])), "gen_s": round(time.time() - t0, 1)}
        A = to_csr_device(dev.edge_index, dev.matrix_values, n, dev.mask)
        b = A.matvec(dev.mask.reshape(-1).double())
        out["b_norm_rel_err"] = abs(float(b.norm()) - float(g["b_norm"])) / float(g["b_norm"])
        cases = [("cuda_G", Gv)]
        if a.oracle_g:
            from oracle import gnn as ognn

            t0 = time.time()
            cpu = sample.to("cpu")
            fwd = ognn.workspace_forward if grid <= 128 else ognn.workspace_forward_chunked
            Go = fwd(sd, cpu.x, cpu.edge_index, cpu.edge_attr, 1).reshape(-1)
            out["oracle_g_s"] = round(time.time() - t0, 1)
            out["oracle_dG_samples"] = float(np.abs(Go.numpy()[g["g_sample_ids"]] - g["g_samples"]).max())
            cases.append(("reference_G", Go.cuda()))
        for name, vals in cases:
            G = to_csr_device(dev.edge_index, vals.reshape(-1, 1, 1), n, dev.mask)
            for mode in modes:
                s = d.DevicePCG(A, mode=mode)
                s.set_spai(G, float(g["epsilon"]))
                x, res, hist = s.solve(b, "ext_spai", rtol=1e-6, history=True)
                h = hist.cpu().numpy()
                ref = g["res_hist"]
                k = min(len(h), len(ref))
                rel = np.abs(h[:k] - ref[:k]) / ref[:k]
                out[f"{name}_mode{mode}"] = {"iters": int(res.iterations), "delta": int(res.iterations) - int(g["iters_pcg"]),
                                             "ms": round(float(res.solve_ms), 2),
                                             "hist_rel_dev_first6": float(rel[:6].max()),
                                             "hist_rel_dev_at": {str(i): float(rel[i]) for i in (10, 50, 100, 200, 400) if i < k},
                                             "last_ref": [float(v) for v in ref[-3:]], "last": [float(v) for v in h[-3:]],
                                             "atol": 1e-6 * float(g["b_norm"])}
                s.close()
            G.close()
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
