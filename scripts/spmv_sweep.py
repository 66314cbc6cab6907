#!/usr/bin/env python
"""Micro-benchmark of the PCG kernels on one Poisson system (development tool, not the bench contract).

    B200SPAI_STREAM_CFG=k B200SPAI_NO_FP32=0|1 python scripts/spmv_sweep.py --grid 216

Prints one JSON line: plain SpMV time for the vector and the streaming kernel, their agreement, and the
per-kernel times of a PCG iteration (b200spai_pcg_profile_iterations).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from learningsparsepreconditioner4gpu_b200 import systems  # noqa: E402
from learningsparsepreconditioner4gpu_b200.device import DeviceCSR, DevicePCG  # noqa: E402


def timed(fn, reps=20):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=216)
    ap.add_argument("--jitter", type=float, default=0.2)
    args = ap.parse_args()
    s = systems.poisson_sample(args.grid, seed=100, jitter=args.jitter, gen_device="cuda", device="cuda")
    n = s.num_nodes
    A = DeviceCSR.from_edges(s.edge_index, s.matrix_values, n, s.mask)
    g = torch.Generator(device="cuda").manual_seed(0)
    gv = (0.1 + 0.002 * torch.randn(s.num_edges, generator=g, device="cuda")).float()
    G = DeviceCSR.from_edges(s.edge_index, gv, n, s.mask)
    x = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
    y1, y2 = torch.empty_like(x), torch.empty_like(x)
    out = {"grid": args.grid, "n": n, "nnzA": A.nnz, "cfg": os.environ.get("B200SPAI_STREAM_CFG", "0"),
           "no_fp32": os.environ.get("B200SPAI_NO_FP32", "0")}
    by = 12 * A.nnz + 20 * n
    for k, y in ((1, y1), (2, y2)):
        ms = timed(lambda: A.matvec(x, out=y, kernel=k))
        out[f"spmv_k{k}_ms"] = round(ms, 4)
        out[f"spmv_k{k}_GBs"] = round(by / ms / 1e6, 1)
    out["k1_vs_k2_maxrel"] = float((y1 - y2).abs().max() / y1.abs().max())
    b = A.matvec(s.mask.reshape(-1).double())
    solver = DevicePCG(A)
    solver.set_spai(G, 3e-3)
    kms = solver.profile_iterations(b, "ext_spai", iterations=20)
    out["pcg_kernel_ms"] = [round(v, 4) for v in kms]
    out["pcg_iter_ms"] = round(sum(kms), 4)
    xs, res, _ = solver.solve(b, "ext_spai", rtol=1e-6, max_iter=200)
    out["graph_ms_per_iter"] = round(res.solve_ms / max(res.iterations, 1), 4)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
