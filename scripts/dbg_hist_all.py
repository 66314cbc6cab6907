"""Debug helper: deviation of the CUDA residual history from the oracle's, per golden case and preconditioner."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from conftest import GOLDEN_CASES, golden_csr, load_golden
from learningsparsepreconditioner4gpu_b200 import device as d
from oracle import pcg as opcg
for name in GOLDEN_CASES:
    g = load_golden(name)
    A, G = golden_csr(g, "A"), golden_csr(g, "G")
    eps = float(g["epsilon"]); gt = g["mask"].reshape(-1).astype(np.float64); b = A @ gt
    Ad, Gd = d.DeviceCSR.from_scipy(A), d.DeviceCSR.from_scipy(G)
    solver = d.DevicePCG(Ad, mode=2); solver.set_spai(Gd, eps)
    for kind, key in (("ext_spai", "iters_pcg"), ("none", "iters_cg"), ("ext_spai_scaled", "iters_pcg_scaled")):
        x, res, hist = solver.solve(torch.from_numpy(b).cuda(), kind, rtol=1e-6, history=True)
        _, its, oh = opcg.pcg_restated(A, b, opcg.make_preconditioner(kind, A, G, eps), rtol=1e-6)
        h = hist.cpu().numpy(); oh = np.asarray(oh); k = min(len(h), len(oh))
        rel = np.abs(h[:k] - oh[:k]) / oh[:k]
        first = lambda t: int(np.argmax(rel > t)) if (rel > t).any() else -1
        print(f"{name:18s} {kind:16s} n={A.shape[0]:5d} it={res.iterations:4d} ref={int(g[key]):4d} oracle={its:4d} "
              f"rel@5={rel[:6].max():.1e} rel@20={rel[:21].max():.1e} first>1e-6:{first(1e-6)} >1e-3:{first(1e-3)} >0.1:{first(0.1)} max={rel.max():.2e}")
        from test_pcg_gpu import history_within_envelope as env
        if its < A.shape[0]:
            print("      envelope violations: (1,.1)", env(h, oh, 1, 0.10), "(2,.25)", env(h, oh, 2, 0.25), "(3,.3)", env(h, oh, 3, 0.3))
