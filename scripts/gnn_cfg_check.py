"""|dG| / max|G| of the GNN kernels selected by B200SPAI_GNN_CFG (2 = FFMA + rational-erfc GELU, 6 = tensor-core 3xTF32
per-edge MLPs, 0 / 4 = the same with erff) against the golden fixtures (reference forward) and the CPU oracle at 40^3, plus
the iteration count of workspace.solve on the CUDA G.  Run once per configuration (the switch is read once per process)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from conftest import GOLDEN_CASES, golden_state_dict, load_golden
from test_gnn_gpu import _sample, _workspace
from learningsparsepreconditioner4gpu_b200 import systems
from oracle import gnn as ognn

cfg = os.environ.get("B200SPAI_GNN_CFG", "default")
for name in GOLDEN_CASES:
    g = load_golden(name)
    ws, s = _workspace(g), _sample(g)
    G = ws(s.x, s.edge_index, s.edge_attr)
    err = np.abs(G.cpu().numpy() - g["G"]).max() / np.abs(g["G"]).max()
    x, res, *_ = ws.solve(s, rtol=1e-6)
    print(f"[cfg {cfg}] golden {name}: dG/max|G| = {err:.2e}  iterations {res.iterations} (reference {int(g['iters_pcg']) if 'iters_pcg' in g else '?'})", flush=True)
for grid in (40, 64):
    sample = systems.poisson_sample(grid, seed=100, gen_device="cpu", device="cpu")
    ws = systems.default_workspace(sample)
    Go = ognn.workspace_forward(ws.state_dict(), sample.x, sample.edge_index, sample.edge_attr, 1)
    dev = sample.to("cuda")
    G = ws(dev.x, dev.edge_index, dev.edge_attr)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): G = ws(dev.x, dev.edge_index, dev.edge_attr)
    torch.cuda.synchronize(); ms = (time.perf_counter() - t0) * 200
    err = float((G.cpu() - Go).abs().max() / Go.abs().max())
    x, res, *_ = ws.solve(dev, rtol=1e-6)
    print(f"[cfg {cfg}] poisson {grid}^3 (E = {sample.num_edges}): dG/max|G| = {err:.2e}, forward {ms:.2f} ms, iterations {res.iterations}", flush=True)
