"""Where is "the reference's iteration count" defined at 216^3?  (CPU only, ~1 h on 8 cores, ~35 GB)

    python scripts/c5_216_band.py [grid=216] > profiles/r02_c5_216_band.jsonl

The committed fixture tests/golden/c5_216.npz holds ONE count of the restated reference recurrence (1451); the CUDA
drivers land at 1378-1403.  This script measures how far the reference's OWN count moves at that size under changes
that are below anything the reference specifies:

  * the summation order of its three dot products (numpy as is / reversed vectors / 64 fixed chunks / long double),
    everything else -- SpMVs, axpys, G -- identical (the study of scripts/cg_order_sensitivity.py at the headline size);
  * G moved by fp32-rounding-level noise: G * (1 + d), d uniform in +-2^-23 (one fp32 ulp, relative), and an absolute
    perturbation uniform in +-2e-6 max|G| (half the measured |dG| of the tensor-core GNN against the oracle),
    with numpy's dots.

The system, weights and G are those of the fixture (seed 100, oracle forward chunked).  One process per variant, forked
after the shared matrices exist.  Each line of the output is one variant: iterations, final recurrence residual, true
residual, max relative history deviation from the fixture, first iteration that deviates by more than 1 %."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def cg(A, G, GT, eps, b, dot, rtol=1e-6):
    """scipy.sparse.linalg.cg's recurrence (oracle.pcg.pcg_restated) with the dot product as a parameter."""
    n = A.shape[0]
    atol = rtol * np.sqrt(dot(b, b))
    x = np.zeros(n)
    r = b.copy()
    hist = [np.sqrt(dot(r, r))]
    rho_prev, p = None, None
    for k in range(n):
        if hist[-1] < atol:
            break
        z = G @ (GT @ r) + eps * r
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
    return x, np.array(hist)


DOTS = {
    "numpy": lambda a, b: float(a @ b),
    "reversed": lambda a, b: float(a[::-1] @ b[::-1]),
    "chunks64": lambda a, b: float(np.sum([float(u @ v) for u, v in zip(np.array_split(a, 64), np.array_split(b, 64))])),
    "longdouble": lambda a, b: float(np.dot(a.astype(np.longdouble), b.astype(np.longdouble))),
}


def main():
    from scipy.sparse import csr_matrix
    from learningsparsepreconditioner4gpu_b200 import systems
    from oracle import assemble as oasm, gnn as ognn

    grid = int(sys.argv[1]) if len(sys.argv) > 1 else 216
    fx = np.load(os.path.join(ROOT, "tests", "golden", f"c5_{grid}.npz"), allow_pickle=False)
    t0 = time.time()
    s = systems.poisson_sample(grid, seed=int(fx["seed"]), gen_device="cpu", device="cpu")
    n, eps = s.num_nodes, float(fx["epsilon"])
    sd = {str(k): torch.from_numpy(fx["sd/" + str(k)].copy()) for k in fx["sd_keys"]}
    cache = f"/tmp/c5_{grid}_oracle_g.npy"
    if os.path.exists(cache):
        g32 = np.load(cache)
    else:
        g32 = ognn.workspace_forward_chunked(sd, s.x, s.edge_index, s.edge_attr, 1).reshape(-1).numpy()
        np.save(cache, g32)
    dev = float(np.abs(g32[fx["g_sample_ids"]] - fx["g_samples"]).max())
    A = oasm.to_csr_cpu_sorted(s.edge_index, s.matrix_values, n, s.mask)
    b = A @ s.mask.numpy().reshape(-1).astype(np.float64)
    ref_hist = fx["res_hist"]
    gmax = float(np.abs(g32).max())
    print(json.dumps({"grid": grid, "n": n, "fixture_iters": int(fx["iters_pcg"]), "setup_s": round(time.time() - t0, 1),
                      "g_vs_fixture_samples_max_abs": dev, "g_max_abs": gmax}), flush=True)

    def g_variant(kind):
        if kind == "oracle":
            v = g32
        else:
            rng = np.random.default_rng({"ulp_rel_a": 1, "ulp_rel_b": 2, "abs2e-6_a": 3, "abs2e-6_b": 4}[kind])
            if kind.startswith("ulp_rel"):
                v = (g32.astype(np.float64) * (1.0 + rng.uniform(-2.0 ** -23, 2.0 ** -23, g32.size))).astype(np.float32)
            else:
                v = (g32.astype(np.float64) + rng.uniform(-2e-6 * gmax, 2e-6 * gmax, g32.size)).astype(np.float32)
        G = oasm.to_csr_cpu_sorted(s.edge_index, torch.from_numpy(v), n, s.mask)
        GT = csr_matrix(G.T)
        GT.sort_indices()
        return G, GT, float(np.abs(v.astype(np.float64) - g32).max() / gmax)

    variants = [("oracle", d) for d in ("numpy", "reversed", "chunks64", "longdouble")] + \
               [(g, "numpy") for g in ("ulp_rel_a", "abs2e-6_a", "abs2e-6_b")]
    only = os.environ.get("BAND_ONLY")
    if only:
        variants = [v for v in variants if f"{v[0]}/{v[1]}" in only.split(",")]
    shared = g_variant("oracle")
    pids = []
    for gk, dk in variants:
        pid = os.fork()
        if pid:
            pids.append(pid)
            continue
        torch.set_num_threads(1)
        t0 = time.time()
        G, GT, dg = shared if gk == "oracle" else g_variant(gk)
        x, h = cg(A, G, GT, eps, b, DOTS[dk])
        k = min(len(h), len(ref_hist))
        rel = np.abs(h[:k] - ref_hist[:k]) / ref_hist[:k]
        out = {"g": gk, "dot": dk, "dg_over_gmax": dg, "iters": len(h) - 1, "final_rel": float(h[-1] / h[0]),
               "true_rel": float(np.linalg.norm(b - A @ x) / np.linalg.norm(b)),
               "max_rel_hist_dev": float(rel.max()), "first_dev_over_1pct": int(np.argmax(rel > 0.01)) if (rel > 0.01).any() else -1,
               "s": round(time.time() - t0, 1)}
        sys.stdout.write(json.dumps(out) + "\n")
        sys.stdout.flush()
        os._exit(0)
    for pid in pids:
        os.waitpid(pid, 0)


if __name__ == "__main__":
    main()
