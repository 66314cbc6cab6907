"""Where the milliseconds of ONE collated batch go (batch.CollatedBatchSolver, BASELINE config C2 mix): phase times with
torch.cuda.synchronize() around each phase, third repetition."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from learningsparsepreconditioner4gpu_b200 import systems
from learningsparsepreconditioner4gpu_b200.device import DevicePCG
from learningsparsepreconditioner4gpu_b200.neural_cg.data import collate
from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

grids = [16, 20, 25, 32]
samples = [systems.poisson_sample(grids[k % 4], seed=1000 + k, device="cuda") for k in range(64)]
ws = systems.default_workspace(samples[0])
for rep in range(3):
    t = {}
    def mark(name, t0):
        torch.cuda.synchronize(); t[name] = (time.perf_counter() - t0) * 1e3; return time.perf_counter()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    batch = collate(samples); t0 = mark("collate", t0)
    n = batch.x.shape[0]
    A = to_csr_device(batch.edge_index, batch.matrix_values, n, batch.mask); t0 = mark("assemble A", t0)
    boo = ws(batch.x, batch.edge_index, batch.edge_attr); t0 = mark("graph + gnn", t0)
    G = to_csr_device(batch.edge_index, boo, n, batch.mask); t0 = mark("assemble G", t0)
    rhs = A.matvec(batch.mask.reshape(-1).double()); t0 = mark("rhs", t0)
    solver = DevicePCG(A); solver.set_spai(G, ws.epsilon); t0 = mark("pcg_create + G^T", t0)
    ptr = batch.ptr.cpu().numpy()
    x, results, _ = solver.solve_batch(ptr, rhs, "ext_spai", rtol=1e-6); t0 = mark("solve_batch", t0)
    its = [r.iterations for r in results]
print({k: round(v, 2) for k, v in t.items()}, "total", round(sum(t.values()), 2), "N", n, "E", batch.edge_index.shape[1],
      "iterations mean/max", np.mean(its), max(its), "solve_ms", results[0].solve_ms)
