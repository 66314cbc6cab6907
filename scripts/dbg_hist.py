"""Debug helper: print the CUDA residual history next to the oracle's for one golden case."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from conftest import golden_csr, load_golden
from learningsparsepreconditioner4gpu_b200 import device as d
from oracle import pcg as opcg
name = sys.argv[1] if len(sys.argv) > 1 else "poisson_zeros_b1"
g = load_golden(name)
A, G = golden_csr(g, "A"), golden_csr(g, "G")
eps = float(g["epsilon"]); gt = g["mask"].reshape(-1).astype(np.float64); b = A @ gt
Ad, Gd = d.DeviceCSR.from_scipy(A), d.DeviceCSR.from_scipy(G)
for mode in (1, 2):
    solver = d.DevicePCG(Ad, mode=mode); solver.set_spai(Gd, eps)
    x, res, hist = solver.solve(torch.from_numpy(b).cuda(), "ext_spai", rtol=1e-6, history=True)
    _, its, oh = opcg.pcg_restated(A, b, opcg.make_preconditioner("ext_spai", A, G, eps), rtol=1e-6)
    h = hist.cpu().numpy()
    print(name, "mode", mode, "iters", res.iterations, "oracle", its, "golden", int(g["iters_pcg"]), "n", A.shape[0])
    for k in range(max(len(h), len(oh))):
        a = h[k] if k < len(h) else float("nan"); o = oh[k] if k < len(oh) else float("nan")
        print(f"  {k:3d} {a:.6e} {o:.6e} {a/o if o else 0:.4f}")
