"""Worker of the multi-GPU row-partition test: run under torchrun, one rank per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 \
        tests/dist_worker.py [grid] [mode]

Every rank builds the same Poisson system, G from the GNN, solves it row-partitioned, and rank 0 also solves it on one
GPU; prints one JSON line with the comparison (asserts on failure)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from learningsparsepreconditioner4gpu_b200 import parallel as par  # noqa: E402
from learningsparsepreconditioner4gpu_b200 import systems  # noqa: E402
from learningsparsepreconditioner4gpu_b200.device import DevicePCG  # noqa: E402
from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device  # noqa: E402


def main():
    grid = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sample = systems.poisson_sample(grid, seed=3, gen_device="cuda", device="cuda")
    ws = systems.default_workspace(sample)
    n = sample.num_nodes
    A = to_csr_device(sample.edge_index, sample.matrix_values, n, sample.mask)
    G = ws.inference_step_device(sample)
    rhs = A.matvec(sample.mask.reshape(-1).double())
    starts = par.partition_rows(par.row_weights_from_edges(sample.edge_index, n), world)
    rs, re = starts[rank], starts[rank + 1]
    ext = par.extended_range(sample.edge_index, rs, re)
    out = {"rank": rank, "world": world, "grid": grid, "rows": [rs, re], "ext": list(ext)}
    for precond in ("ext_spai", "none", "diagonal"):
        spai = precond == "ext_spai"
        solver = par.DistPCG(A, G if spai else None, ws.epsilon, starts, ext, mode=mode if (spai or mode != 4) else 0, reuse=True)
        assert solver.mode == mode or not spai
        x_loc, res, hist = solver.solve(rhs[rs:re].contiguous(), precond, rtol=1e-8, history=True)
        x = solver.gather(x_loc)
        iters = [None] * world
        dist.all_gather_object(iters, int(res.iterations))
        assert len(set(iters)) == 1, f"ranks disagree on the iteration count: {iters}"
        assert res.converged == 1 and res.breakdown == 0, (res.converged, res.breakdown)
        # reference: the single-GPU solver on the same matrices
        ref = DevicePCG(A)
        if precond == "ext_spai":
            ref.set_spai(G, ws.epsilon)
        x1, res1, hist1 = ref.solve(rhs, precond, rtol=1e-8, history=True)
        assert res.reserved == (mode or 2) or not spai, (res.reserved, mode)
        assert abs(res.iterations - res1.iterations) <= 1, (precond, res.iterations, res1.iterations)
        k = min(6, len(hist), len(hist1))
        assert torch.allclose(hist[:k], hist1[:k], rtol=1e-10, atol=0), (hist[:k], hist1[:k])
        rel = float((x - x1).norm() / x1.norm())
        assert rel < 1e-6, (precond, rel)
        true_rel = float((rhs - A.matvec(x)).norm() / rhs.norm())
        assert true_rel < 1e-7, true_rel
        out[precond] = {"iters": int(res.iterations), "iters_1gpu": int(res1.iterations), "x_rel_diff": rel,
                        "true_rel_res": true_rel, "ms": float(res.solve_ms), "ms_1gpu": float(res1.solve_ms), **solver.info()}
        solver.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    par.DistPCG.drop_connections()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
