"""How far does the REFERENCE's own iteration count move when only the summation order of its dot products changes?

    python scripts/cg_order_sensitivity.py 64 128 heat elast synth10240 synth15360 synth20480

The scipy cg recurrence (oracle.pcg.pcg_restated, pinned to the reference's get_pcg_iter_time_scipy by the C5 fixtures)
is run on the fixture systems with the oracle's G, four times: numpy's dots as they are, dots over the reversed vectors,
dots accumulated in long double, and dots summed in 64 fixed chunks (a GPU-like tree).  SpMVs, axpys and the
preconditioner are identical in all four runs -- every difference comes from the last bits of rho, p.q and ||r||.
The spread of the four counts is the band inside which "the same iteration count" is defined for this system; it is
the tolerance tests/test_c5_gpu.py states.  CPU only; results are recorded in DESIGN.md §4."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def cg(A, b, M, dot, rtol=1e-6):
    n = A.shape[0]
    atol = rtol * np.sqrt(dot(b, b))
    x = np.zeros(n)
    r = b.copy()
    hist = [np.sqrt(dot(r, r))]
    rho_prev, p = None, None
    for k in range(n):
        if hist[-1] < atol:
            break
        z = M(r)
        rho = dot(r, z)
        if k > 0:
            p *= rho / rho_prev
            p += z
        else:
            p = z.copy()
        q = A @ p
        alpha = rho / dot(p, q)
        x += alpha * p
        r -= alpha * q
        rho_prev = rho
        hist.append(np.sqrt(dot(r, r)))
    return len(hist) - 1, np.array(hist)


def main():
    from learningsparsepreconditioner4gpu_b200 import systems
    from oracle import assemble as oasm, gnn as ognn, pcg as opcg

    dots = {
        "numpy": lambda a, b: float(a @ b),
        "reversed": lambda a, b: float(a[::-1] @ b[::-1]),
        "longdouble": lambda a, b: float(np.dot(a.astype(np.longdouble), b.astype(np.longdouble))),
        "chunks64": lambda a, b: float(np.sum([float(u @ v) for u, v in zip(np.array_split(a, 64), np.array_split(b, 64))])),
    }
    for tok in sys.argv[1:]:
        if tok.isdigit():        # C5 family: the committed fixture pins the reference's count and history
            grid = int(tok)
            g = np.load(os.path.join(ROOT, "tests", "golden", f"c5_{grid}.npz"), allow_pickle=False)
            s = systems.poisson_sample(grid, seed=int(g["seed"]), gen_device="cpu", device="cpu")
            sd = {str(k): torch.from_numpy(g["sd/" + str(k)].copy()) for k in g["sd_keys"]}
            G = ognn.workspace_forward(sd, s.x, s.edge_index, s.edge_attr, 1)
            n, eps = s.num_nodes, float(g["epsilon"])
            A = oasm.to_csr_cpu_sorted(s.edge_index, s.matrix_values, n, s.mask)
            Gc = oasm.to_csr_cpu_sorted(s.edge_index, G.reshape(-1), n, s.mask)
            out = {"grid": grid, "reference_iters": int(g["iters_pcg"])}
            ref_hist = g["res_hist"]
        else:                    # C1 / C3 / C4 at their BASELINE sizes (systems.shape_sample): the oracle is the reference
            s, bs = systems.shape_sample(tok)
            ws = systems.default_workspace(s, block_size=bs)
            G = ognn.workspace_forward(ws.state_dict(), s.x, s.edge_index, s.edge_attr, bs)
            n, eps = s.num_nodes * bs, ws.epsilon
            A = oasm.to_csr_cpu(s.edge_index, s.matrix_values, n, s.mask)
            Gc = oasm.to_csr_cpu(s.edge_index, G, n, s.mask)
            out, ref_hist = {"shape": tok}, None
        b = A @ s.mask.numpy().reshape(-1).astype(np.float64)
        M = opcg.make_preconditioner("ext_spai", A, Gc, eps)
        for name, dot in dots.items():
            t0 = time.time()
            it, h = cg(A, b, M, dot)
            if ref_hist is None:
                ref_hist = h     # the "numpy" run is the reference recurrence itself
                out["reference_iters"] = it
            k = min(len(h), len(ref_hist))
            rel = np.abs(h[:k] - ref_hist[:k]) / ref_hist[:k]
            out[name] = {"iters": it, "max_rel_hist_dev": float(rel.max()), "first_dev_over_1pct": int(np.argmax(rel > 0.01)) if (rel > 0.01).any() else -1,
                         "s": round(time.time() - t0, 1)}
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
