"""Worker of the multi-GPU partitioned-system test: run under torchrun, one rank per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29631 \
        tests/dist_gnn_worker.py [poisson|elast] [grid]

Every rank holds the inputs of the same system but computes only its share: GNN on the edges whose target it owns
(node-feature halo refreshed after every layer), boundary G swap, assembly of its rows of A, G, G^T, row-partitioned
PCG.  For the check each rank ALSO runs the single-GPU path on its own device and compares bit for bit (G values,
matrices) and to +-1 iteration (solve).  Prints one JSON line on rank 0 (asserts on failure)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from learningsparsepreconditioner4gpu_b200 import parallel as par  # noqa: E402
from learningsparsepreconditioner4gpu_b200 import systems  # noqa: E402
from learningsparsepreconditioner4gpu_b200.device import DevicePCG  # noqa: E402
from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device  # noqa: E402


def same(a, b):
    a, b = a.to_scipy(), b
    return (a.shape == b.shape and np.array_equal(a.indptr, b.indptr) and np.array_equal(a.indices, b.indices)
            and np.array_equal(a.data, b.data))


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "poisson"
    grid = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if kind == "elast":
        sample, b = systems.elasticity_sample(grid, 6, 6, seed=1, device="cuda"), 3
    else:
        sample, b = systems.poisson_sample(grid, seed=3, gen_device="cuda", device="cuda"), 1
    ws = systems.default_workspace(sample, block_size=b)
    N = sample.num_nodes
    n = N * b
    starts = par.partition_rows(par.row_weights_from_edges(sample.edge_index, N), world)
    ps = par.PartitionedSystem(sample.edge_index, N, b, starts)
    g_t = ps.gnn_forward(ws.gnn, sample.x, sample.edge_attr)
    A_loc, G_loc, GT_loc = ps.assemble(sample.matrix_values, g_t, sample.mask)
    # single-GPU reference on this rank's device
    g_ref = ws(sample.x, sample.edge_index, sample.edge_attr)                 # [E, b, b]
    assert torch.equal(g_t.reshape(-1, b, b), g_ref[ps.ids_t]), "partitioned GNN differs from the single-GPU forward"
    assert torch.equal(ps.g_rows(g_t.reshape(-1, b, b)), g_ref[ps.e0:ps.e1]), "boundary G swap"
    A = to_csr_device(sample.edge_index, sample.matrix_values, n, sample.mask)
    G = to_csr_device(sample.edge_index, g_ref, n, sample.mask)
    rs, re, shift, hi = ps.layout
    lo = max(shift, 0)

    def window(M):  # global scipy rows [rs, re), columns [shift, hi) with the (unused) slot of a negative shift in front
        import scipy.sparse as sp
        W = M[rs:re, lo:hi]
        return sp.csr_matrix(sp.hstack([sp.csr_matrix((re - rs, lo - shift)), W])) if lo > shift else sp.csr_matrix(W)

    As, Gs = A.to_scipy(), G.to_scipy()
    for name, loc, glob in (("A", A_loc, As), ("G", G_loc, Gs), ("GT", GT_loc, Gs.T.tocsr())):
        w = window(glob)
        w.sort_indices()
        assert same(loc, w), f"local rows of {name} differ from the global matrix"
    mask64 = sample.mask.reshape(-1).double()
    rhs = A.matvec(mask64)
    rhs_loc = A_loc.matvec(ps.ext_slice(sample.mask.reshape(N, b)).reshape(-1).double())
    assert torch.equal(rhs_loc, rhs[rs:re])
    solver = ps.make_solver((A_loc, G_loc, GT_loc), ws.epsilon)
    x_loc, res, hist = solver.solve(rhs_loc, "ext_spai", rtol=1e-8, history=True)
    x = solver.gather(x_loc)
    ref = DevicePCG(A)
    ref.set_spai(G, ws.epsilon)
    x1, res1, hist1 = ref.solve(rhs, "ext_spai", rtol=1e-8, history=True)
    assert res.converged == 1 and abs(res.iterations - res1.iterations) <= 1, (res.iterations, res1.iterations)
    rel = float((x - x1).norm() / x1.norm())
    assert rel < 1e-6, rel
    out = {"rank": rank, "world": world, "kind": kind, "N": N, "E": sample.num_edges, "edges_gnn": int(ps.ids_t.numel()),
           "sent": int(sum(v.numel() for v in ps.plan[5].values())), "iters": int(res.iterations), "iters_1gpu": int(res1.iterations),
           "x_rel_diff": rel}
    dist.barrier()
    solver.close()
    ps.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
