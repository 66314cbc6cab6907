"""Host-side phase times of one small solve (where the per-system milliseconds go)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from learningsparsepreconditioner4gpu_b200 import systems
from learningsparsepreconditioner4gpu_b200.device import DevicePCG
from learningsparsepreconditioner4gpu_b200.neural_cg.nn.gnns import DeviceGraph
from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

g = int(sys.argv[1]) if len(sys.argv) > 1 else 20
s = systems.poisson_sample(g, seed=1, device="cuda")
ws = systems.default_workspace(s)
n = s.num_nodes


def T():
    torch.cuda.synchronize()
    return time.perf_counter()


for mode in (3, 2):
    acc = {}
    for rep in range(12):
        t = [T()]
        graph = DeviceGraph(s.edge_index, n); t.append(T())
        _, gv = ws.gnn(s.x, s.edge_index, s.edge_attr, graph=graph); t.append(T())
        A = to_csr_device(s.edge_index, s.matrix_values, n, s.mask); t.append(T())
        G = to_csr_device(s.edge_index, gv.reshape(-1, 1, 1), n, s.mask); t.append(T())
        rhs = A.matvec(s.mask.reshape(-1).double()); t.append(T())
        solver = DevicePCG(A, mode=mode); t.append(T())
        solver.set_spai(G, ws.epsilon); t.append(T())
        x, res, _ = solver.solve(rhs, "ext_spai", rtol=1e-6); t.append(T())
        x2, res2, _ = solver.solve(rhs, "ext_spai", rtol=1e-6); t.append(T())
        del solver, A, G; graph.close(); t.append(T())
        if rep >= 2:
            for name, a, b in zip(["graph_create", "gnn", "A csr", "G csr", "rhs", "pcg_create", "set_spai", "solve#1", "solve#2", "free"], t[:-1], t[1:]):
                acc[name] = acc.get(name, 0.0) + (b - a) * 1e3 / 10
    print(f"grid {g} mode {mode} iters {res.iterations} device solve_ms {res.solve_ms:.3f}:", {k: round(v, 3) for k, v in acc.items()}, "total", round(sum(acc.values()), 2))
