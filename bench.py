#!/usr/bin/env python
"""bench.py -- the learned-SPAI solve path on B200: GNN -> G -> CSR assembly -> PCG to rtol 1e-6.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--grid 216] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json north_star target / config C5 sized for ONE GPU): a 3-D Poisson system on a jittered
Kuhn-tet grid of grid^3 nodes (216^3 = 10.08 M DOF, ~15 stored entries per row), Dirichlet patch, random-init
GNN weights (seed 42), rhs b = A * mask, x0 = 0, rtol 1e-6, fp64 solve.  One STEP = one system through the whole
path: topology preprocessing, GNN forward (G values), assembly of A and G into masked fp64 CSR, G^T, PCG.
With N GPUs every rank solves its own independent system of the same shape (weak scaling, no data-path
collective -- the reference treats batch members as independent systems, workspace.py:137-143).

One JSON line on stdout (rank 0):  value = solves/s with the inputs resident in HBM; e2e = the same through
``workspace.solve`` with the sample in pinned HOST memory (H2D of every input, D2H of x inside the timed region);
roofline = the PCG SpMV kernel family against MEASURED_PEAKS.json; cpu_baseline = the CPU oracle (scipy PCG +
torch GNN restatement of the reference) timed on a bounded sample on this box.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

FUSED_PHASES = ("GT_r+barrier", "G_t+barrier", "update_p+barrier", "A_p+barrier", "update_xr+barrier")
METRIC = "learned-SPAI PCG solves/s to rtol 1e-6 (GNN->G, assembly, PCG; time-to-1e-6 in extra keys)"
UNIT = "solves/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ---------------------------------------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------------------------------------
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                f = [t.strip() for t in line.split(",")]
                if len(f) < 7:
                    continue
                try:
                    sm.append(float(f[0]))
                    mx.append(float(f[1]))
                except ValueError:
                    continue
                for nm, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def ncu_traffic(grid: int):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the streaming SpMV from the newest committed ncu summary
    (profiles/r*_spmv_stream_ncu_full.txt, 216^3 only) -- (bytes or None, source)."""
    import glob
    import re

    if grid != 216:
        return None, "no ncu capture for this grid size"
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_spmv_stream_ncu_full.txt")))
    if not files:
        return None, "no ncu summary under profiles/"
    rd, wr = [], []
    for line in open(files[-1]):
        m = re.match(r"dram__bytes_(read|write)\.sum \[(\w+)\]:\s*\[(.*)\]", line.strip())
        if m:
            scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[m.group(2)]
            vals = [float(v.strip().strip("'")) * scale for v in m.group(3).split(",")]
            (rd if m.group(1) == "read" else wr).extend(vals)
    if not rd or len(rd) != len(wr):
        return None, f"could not parse {os.path.basename(files[-1])}"
    return float(np.mean([a + b for a, b in zip(rd, wr)])), f"ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, mean of {len(rd)} launches ({os.path.relpath(files[-1], ROOT)})"


def workload_text(g, N, E, rtol) -> str:
    """The ONE description of the workload both arms (and every N) print, so that their `config` dicts are equal."""
    return (f"poisson3d jittered Kuhn-tet grid {g}^3 = {N} DOF, E = {E} stored entries, Dirichlet patch, "
            f"random-init GNN (seed 42), eps 3e-3, rhs = A*mask, rtol {rtol:g}")


def tensor_bytes(*ts) -> int:
    return int(sum(t.numel() * t.element_size() for t in ts if isinstance(t, torch.Tensor)))


INPUT_FIELDS = ("x", "edge_index", "edge_attr", "mask", "matrix_values")


# ---------------------------------------------------------------------------------------------------------
# the CUDA arm
# ---------------------------------------------------------------------------------------------------------
def run_step(ws, sample, rtol, timers=None):
    """One system through the whole path on the current device.  ``sample`` tensors are already on the GPU."""
    from learningsparsepreconditioner4gpu_b200.device import DevicePCG
    from learningsparsepreconditioner4gpu_b200.neural_cg.nn.gnns import DeviceGraph
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    def mark(name):
        if timers is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            timers.append((name, ev))

    n = sample.x.shape[0] * ws.block_size
    mark("start")
    graph = DeviceGraph(sample.edge_index, sample.x.shape[0])           # target-sorted edges, segment offsets
    mark("graph")
    _, g_vals = ws.gnn(sample.x, sample.edge_index, sample.edge_attr, graph=graph)   # G values, fp32 [E, b*b]
    g_vals = g_vals.reshape(-1, ws.block_size, ws.block_size)
    mark("gnn")
    A = to_csr_device(sample.edge_index, sample.matrix_values, n, sample.mask)       # masked fp64 CSR
    G = to_csr_device(sample.edge_index, g_vals, n, sample.mask)
    rhs = A.matvec(sample.mask.reshape(-1).to(torch.float64))                         # b = A @ mask (infer.py:297)
    mark("assemble")
    solver = DevicePCG(A)
    solver.set_spai(G, ws.epsilon)                                                    # G^T on the device
    mark("transpose")
    x, res, _ = solver.solve(rhs, "ext_spai", rtol=rtol)
    mark("pcg")
    graph.close()
    return x, res, solver, A, G, rhs


def phases_ms(timers):
    torch.cuda.synchronize()
    out = {}
    for (_, e0), (name, e1) in zip(timers[:-1], timers[1:]):
        out[name] = out.get(name, 0.0) + e0.elapsed_time(e1)
    return out


class Ctx:
    """One process per GPU: rank / world from the torchrun environment, NCCL process group as plumbing (barriers,
    max-over-ranks of the device times)."""

    def __init__(self):
        from learningsparsepreconditioner4gpu_b200 import _lib

        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        _lib.require_cuda()
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist

            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29533")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local), rank=self.rank, world_size=self.world)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def _reduce(self, v: float, op) -> float:
        if self.dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, v: float) -> float:
        return self._reduce(v, None if self.dist is None else self.dist.ReduceOp.MAX)

    def sum(self, v: float) -> float:
        return self._reduce(v, None if self.dist is None else self.dist.ReduceOp.SUM)

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def main_ours(args, ctx: Ctx, emit: bool = True):
    """Independent systems, one per GPU (weak scaling; the whole bench line at N = 1, the "replicas" figure at N > 1)."""
    from learningsparsepreconditioner4gpu_b200 import _lib, systems

    rank, world, local, dist = ctx.rank, ctx.world, ctx.local, ctx.dist
    n_gpus = world
    if args.gpus and args.gpus != world and rank == 0:
        log(f"[bench] --gpus {args.gpus} but WORLD_SIZE={world}: running on {world} rank(s)")
    barrier, max_over_ranks, sum_over_ranks = ctx.barrier, ctx.max, ctx.sum

    g = args.grid
    t0 = time.time()
    sample = systems.poisson_sample(g, seed=100 + rank, gen_device="cuda", device="cuda")
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    N, E = sample.num_nodes, sample.num_edges
    ws = systems.default_workspace(sample)
    host = sample.to("cpu").pin_memory()                      # the e2e arm starts from pinned host memory
    log(f"[bench r{rank}] system {g}^3: N={N} E={E} generated in {time.time() - t0:.1f}s")
    h2d = tensor_bytes(*[getattr(host, k) for k in INPUT_FIELDS])
    d2h = N * ws.block_size * 8

    # ---- warm-up (also builds the lazily created pieces: weight blob, occupancy queries, graphs) ----------
    for _ in range(max(args.warmup, 3)):
        x, res, solver, A, G, rhs = run_step(ws, sample, args.rtol)
        del solver, A, G, rhs
    iters_warm = res.iterations
    log(f"[bench r{rank}] warm-up done: {res.iterations} iterations, converged={res.converged}, pcg {res.solve_ms:.1f} ms")

    # ---- timed region 1: inputs resident in HBM -------------------------------------------------------------
    sampler = ClockSampler(local)
    launches0 = _lib.total_launches()
    timers = []
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    pcg_ms, iters = [], []
    for _ in range(args.steps):
        x, res, solver, A, G, rhs = run_step(ws, sample, args.rtol, timers)
        pcg_ms.append(float(res.solve_ms))
        iters.append(int(res.iterations))
        if _ < args.steps - 1:
            del solver, A, G, rhs
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = _lib.total_launches() - launches0
    ph = phases_ms(timers)
    converged = int(res.converged)
    value = n_gpus * args.steps / (dev_ms * 1e-3)

    pcg_mode = int(res.reserved)
    fused_phases = None
    if pcg_mode == 4:
        fp_ms, fp_it = solver.fused_profile()
        fused_phases = {"iterations": fp_it, "ms_per_iteration": {k: round(v / max(fp_it, 1), 4) for k, v in
                                                                  zip(FUSED_PHASES, fp_ms)}}
    # ---- per-kernel times of the PCG iteration (roofline), on the last step's matrices ----------------------
    kms = solver.profile_iterations(rhs, "ext_spai", iterations=20)
    nnzA, nnzG, n = A.nnz, G.nnz, N * ws.block_size
    bytes_k = {
        "GT_r (t = G^T r)": 12 * nnzG + 20 * n + 4,
        "G_t (z = G t + eps r; r.z)": 12 * nnzG + 28 * n + 4,
        "update_p": 24 * n,
        "A_p (q = A p; p.q)": 12 * nnzA + 20 * n + 4,
        "update_xr (x, r; r.r)": 48 * n,
    }
    peak, peak_src = peaks()
    # bytes the kernels really read: lossless fp32 value stream and 16-bit column steps where they apply
    (va, ia), (vg, ig) = A.stream_bytes(), G.stream_bytes()
    per_a, per_g = va + ia, vg + ig
    row_a, row_g = (8 if ia == 2 else 4), (8 if ig == 2 else 4)      # row pointer (+ first column of the row)
    streamed_k = {
        "GT_r (t = G^T r)": per_g * nnzG + (row_g + 16) * n,
        "G_t (z = G t + eps r; r.z)": per_g * nnzG + (row_g + 24) * n,
        "update_p": 24 * n,
        "A_p (q = A p; p.q)": per_a * nnzA + (row_a + 16) * n,
        "update_xr (x, r; r.r)": 48 * n,
    }
    detail = {}
    for (name, by), ms in zip(bytes_k.items(), kms):
        detail[name] = {"ms": round(ms, 4), "bytes": by, "GB/s": round(by / (ms * 1e-3) / 1e9, 1) if ms > 0 else None,
                        "streamed_bytes": streamed_k[name],
                        "streamed_GB/s": round(streamed_k[name] / (ms * 1e-3) / 1e9, 1) if ms > 0 else None}
    spmv_bytes = sum(by for k, by in bytes_k.items() if "update" not in k)
    spmv_streamed = sum(by for k, by in streamed_k.items() if "update" not in k)
    spmv_ms = kms[0] + kms[1] + kms[3]
    algorithmic = spmv_bytes / 3 / (spmv_ms / 3 * 1e-3) / 1e9
    streamed = spmv_streamed / 3 / (spmv_ms / 3 * 1e-3) / 1e9
    iter_bytes = sum(bytes_k.values())
    iter_streamed = sum(streamed_k.values())
    iter_ms = sum(kms)
    vt = {4: "float", 8: "double"}
    ct = {2: "unsigned short (16-bit column steps)", 4: "int"}
    traffic, traffic_src = ncu_traffic(g)
    roofline = {
        "kernel": f"spmv_stream_kernel<4096,2,8,{vt.get(va, va)},{ct.get(ia, ia)},*> (3 launches per PCG iteration: t=G^T r, z=G t+eps r, q=A p)"
                  if nnzA >= (1 << 22) else "spmv kernels of the 5-launch iteration",
        # achieved / frac: the bytes the kernel REALLY streams per launch (lossless fp32 values + 16-bit column steps where they
        # apply; agrees with ncu's dram__bytes, `traffic`) over its CUDA-event time -- a true fraction of the measured HBM peak
        "bound": "hbm", "achieved": round(streamed, 1), "peak": peak, "unit": "GB/s", "frac": round(streamed / peak, 4),
        "peak_source": peak_src, "traffic": traffic, "traffic_source": traffic_src,
        "bytes_per_launch": spmv_streamed // 3, "ms_per_launch": round(spmv_ms / 3, 4),
        # the same time against SURVEY 8d's algorithmic bytes (12 B per entry: fp64 value + i32 index): what a kernel streaming the
        # uncompressed CSR would have to sustain -- > peak means the compression buys more than the roofline allows without it
        "algorithmic": {"bytes_per_launch": spmv_bytes // 3, "GB/s": round(algorithmic, 1), "frac_of_peak": round(algorithmic / peak, 4),
                        "speedup_from_lossless_stream": round(spmv_bytes / max(spmv_streamed, 1), 3)},
        "pcg_iteration": {"ms": round(iter_ms, 4), "streamed_bytes": iter_streamed, "streamed_GB/s": round(iter_streamed / (iter_ms * 1e-3) / 1e9, 1),
                          "frac": round(iter_streamed / (iter_ms * 1e-3) / 1e9 / peak, 4), "algorithmic_bytes": iter_bytes,
                          "algorithmic_GB/s": round(iter_bytes / (iter_ms * 1e-3) / 1e9, 1)},
        "per_kernel": detail,
    }
    # GNN forward against the fp32 FFMA roofline (MAC counts per edge / node; 2 flop per MAC)
    Fe, Fin, bb = sample.edge_attr.shape[1], sample.x.shape[1], ws.block_size ** 2
    gnn_flop = 2.0 * (E * (11776 + 16 * Fe + 16 * bb) + N * (4 * 768 + 16 * Fin + 512))
    gnn_ms = ph.get("gnn", 0.0) / args.steps
    ffma_peak = 148 * 128 * 2 * 1.965e9 / 1e12
    gnn_cfg = os.environ.get("B200SPAI_GNN_CFG", "6")
    gnn_tensor = gnn_cfg.isdigit() and (int(gnn_cfg) & 4) != 0
    roofline["gnn"] = {"kernel": ("gnn_edge_mma_kernel (per-edge MLPs as 3xTF32 mma.sync, fp32 accumulate: 3 tensor MACs per algorithmic MAC) + "
                                  "gnn_node_kernel (fp32 FFMA) x 4 MP layers, encoders, gnn_dec_mma_kernel" if gnn_tensor else
                                  "gnn_edge_kernel + gnn_node_kernel x 4 MP layers, encoders, decoder (fp32 FFMA)"),
                       "bound": "fp32 (algorithmic flops against the FFMA roofline; ncu of the edge kernel: tensor pipe 44 % active, issue 57 %, "
                                "profiles/r02_gnn_edge_mma_ncu_full.txt)" if gnn_tensor else "fp32",
                       "flop": gnn_flop, "ms": round(gnn_ms, 3), "achieved": round(gnn_flop / (gnn_ms * 1e-3) / 1e12, 2) if gnn_ms > 0 else None,
                       "peak": round(ffma_peak, 1), "unit": "TFLOP/s",
                       "frac": round(gnn_flop / (gnn_ms * 1e-3) / 1e12 / ffma_peak, 4) if gnn_ms > 0 else None,
                       "peak_source": "148 SMs x 128 FFMA lanes x 2 flop x 1.965 GHz (SURVEY 8d); frac_at_sampled_clock uses the SM clock under load"}
    if clocks and clocks.get("sm_mhz") and gnn_ms > 0:
        roofline["gnn"]["frac_at_sampled_clock"] = round(gnn_flop / (gnn_ms * 1e-3) / 1e12 / (148 * 128 * 2 * clocks["sm_mhz"] * 1e6 / 1e12), 4)
    del solver, A, G, rhs
    torch.cuda.empty_cache()

    # ---- timed region 2: end to end from pinned host memory -------------------------------------------------
    x_host = torch.empty(N * ws.block_size, dtype=torch.float64).pin_memory()

    def e2e_step():
        dev = host.to("cuda", non_blocking=True)              # H2D of x, edge_index, edge_attr, mask, matrix_values
        xs, r, sv, A_, G_, b_ = run_step(ws, dev, args.rtol)
        x_host.copy_(xs, non_blocking=True)                    # D2H of the solution
        torch.cuda.current_stream().synchronize()
        return r

    e2e_step()
    barrier()
    ev0.record()
    for _ in range(args.steps):
        r2 = e2e_step()
    ev1.record()
    barrier()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1))
    e2e_value = n_gpus * args.steps / (e2e_ms * 1e-3)

    # ---- CPU oracle beside it (rank 0, N = 1 only) -----------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu = cpu_reference_sample(sample, ws, int(np.mean(iters)), args)
        except Exception as exc:  # the baseline must never take the bench line down with it
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {exc!r}"}

    total_iters = sum_over_ranks(float(np.mean(iters))) / n_gpus
    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(dev_ms / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64 (PCG) / f32 (GNN)", "data": "synthetic",
            "config": {"workload": workload_text(g, N, E, args.rtol), "systems_per_step": n_gpus},
            "parallelism": f"{n_gpus} independent system(s), one per GPU",
            "l2": "inputs larger than L2 (A + G + G^T = %.1f GB streamed per iteration)" % ((nnzA * (va + ia) + 2 * nnzG * (vg + ig)) / 1e9),
            "time_to_tol_ms": round(float(np.mean(pcg_ms)), 3), "iterations": round(total_iters, 1), "converged": converged,
            "ms_per_iteration": round(float(np.mean(pcg_ms)) / max(float(np.mean(iters)), 1.0), 4),
            "pcg_mode": pcg_mode, "fused_phases": fused_phases,
            "phases_ms_per_step": {k: round(v / args.steps, 3) for k, v in ph.items()},
            "e2e": {"value": round(e2e_value, 4), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": round(e2e_ms / args.steps, 3), "api": "SimpleTrainingWorkspace.solve path on a pinned-host sample"},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
        }
        if emit:
            print(json.dumps(line), flush=True)
        return line
    return None


# ---------------------------------------------------------------------------------------------------------
# the CPU arm: the oracle (restated reference) on this box's host cores
# ---------------------------------------------------------------------------------------------------------
class CpuArm:
    """Bounded CPU sample of one step of the same workload (the oracle = restated reference), scaled to the
    metric's unit.

    * GNN: the oracle forward (torch CPU, all threads) on a gnn_grid^3 system of the same family; the per-edge time
      is scaled to the workload's E (the forward is O(E)).
    * PCG: ``pcg_iters`` iterations of the oracle's scipy solve (validate.py:163-201: scipy cg, M = G G^T + eps I,
      G^T materialised as CSR before the clock starts) on the FULL system; per-iteration time scaled to
      ``iterations``.  scipy's csr_matvec is single-threaded.  G's values are a value-independent timing stand-in
      on A's pattern (|G| ~ 0.1 like the random-init GNN's output; csr_matvec cost does not depend on values), so
      nothing computed by the CUDA library enters this arm.
    """

    def __init__(self, sample_cpu, ws, rtol, gnn_grid):
        from learningsparsepreconditioner4gpu_b200 import systems
        from oracle import assemble as oasm
        from oracle import pcg as opcg

        self.threads = torch.get_num_threads()
        self.rtol = rtol
        self.small = systems.poisson_sample(gnn_grid, seed=7)
        self.gnn_grid = gnn_grid
        self.sd = {k: v.detach().cpu() for k, v in ws.state_dict().items()}
        self.E = sample_cpu.num_edges
        n = sample_cpu.num_nodes
        A = oasm.to_csr_cpu_sorted(sample_cpu.edge_index, sample_cpu.matrix_values, n, sample_cpu.mask)
        rng = np.random.default_rng(0)
        g_vals = (0.1 + 0.002 * rng.standard_normal(self.E)).astype(np.float32)
        G = oasm.to_csr_cpu_sorted(sample_cpu.edge_index, g_vals, n, sample_cpu.mask)
        self.gt = sample_cpu.mask.numpy().reshape(-1).astype(np.float64)
        self.timer = opcg.ScipyPcgTimer(A, G, ws.epsilon)

    def step(self, iterations: int, pcg_iters: int):
        from oracle import gnn as ognn

        s = self.small
        t0 = time.perf_counter()
        ognn.workspace_forward(self.sd, s.x, s.edge_index, s.edge_attr, 1)
        t_gnn_small = time.perf_counter() - t0
        t_gnn = t_gnn_small * self.E / s.num_edges
        done, t_pcg = self.timer.run(self.gt, self.rtol, pcg_iters)
        t_iter = t_pcg / max(done, 1)
        t_solve = t_iter * iterations
        return {
            "value": 1.0 / (t_gnn + t_solve), "unit": UNIT, "cores": self.threads, "kind": "port",
            "sample": (f"oracle GNN forward on a {self.gnn_grid}^3 system ({s.num_edges} edges, {t_gnn_small:.2f} s, "
                       f"{self.threads} torch threads) scaled by E to {t_gnn:.1f} s; oracle scipy PCG (1 thread) {done} iterations "
                       f"on the full system = {t_iter:.3f} s/iteration x {iterations} iterations = {t_solve:.1f} s; extrapolated"),
            "gnn_s": t_gnn, "pcg_s": t_solve, "s_per_iteration": t_iter, "iterations": iterations,
        }


def cpu_reference_sample(sample, ws, iterations: int, args):
    """cpu_baseline of the CUDA arm: scaled with the oracle's OWN iteration count when a fixture holds it (216^3), else
    with the CUDA path's count."""
    fx = os.path.join(ROOT, "tests", "golden", f"c5_{args.grid}.npz")
    if os.path.exists(fx):
        iterations = int(np.load(fx, allow_pickle=False)["iters_pcg"])
    arm = CpuArm(sample.to("cpu"), ws, args.rtol, gnn_grid=min(96, args.grid))
    arm.step(iterations, 1)  # warm-up (page in, torch thread pool)
    return arm.step(iterations, 6)


def reference_iterations(grid: int):
    """(iterations, source) the reference's solve of the bench system takes.  216^3: the oracle run to completion on the
    CPU-generated twin of the bench system (tests/golden/c5_216.npz, minted by tests/golden/make_golden_c5.py --oracle 216:
    1451 iterations in 1608 s on 4 cores of the build container); other sizes: the committed fixture when there is one."""
    env = os.environ.get("B200SPAI_REF_ITERS")
    if env:
        return int(env), "B200SPAI_REF_ITERS"
    fx = os.path.join(ROOT, "tests", "golden", f"c5_{grid}.npz")
    if os.path.exists(fx):
        return int(np.load(fx, allow_pickle=False)["iters_pcg"]), f"tests/golden/c5_{grid}.npz (oracle solve run to completion)"
    return int(6.7 * grid), "6.7 x grid edge (fit through the 64^3 / 128^3 / 216^3 fixtures)"


def main_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the reference itself needs torch_geometric,
    lightning and pymathprim, none installable here) on this box's host cores, same workload and metric.  Every step is a
    bounded sample: the oracle GNN forward on a 96^3 system of the same family scaled by E, and 3 iterations of the
    oracle's scipy PCG on the FULL system scaled by the iteration count of the oracle's own complete solve (fixture)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from learningsparsepreconditioner4gpu_b200 import systems

    torch.set_num_threads(os.cpu_count() or 1)     # torchrun exports OMP_NUM_THREADS=1: use every host core at every N
    g = args.grid
    iterations, it_src = reference_iterations(g)
    gen_device = "cuda" if torch.cuda.is_available() else "cpu"   # input generation only
    sample = systems.poisson_sample(g, seed=100, gen_device=gen_device, device="cpu")
    ws = systems.default_workspace(sample)
    N, E = sample.num_nodes, sample.num_edges
    arm = CpuArm(sample, ws, args.rtol, gnn_grid=min(96, g))
    steps = max(args.steps, 1)
    vals, last = [], None
    t_begin = time.perf_counter()
    for i in range(max(args.warmup, 0) + steps):
        last = arm.step(iterations, pcg_iters=3)
        if i >= args.warmup:
            vals.append(last["value"])
        if time.perf_counter() - t_begin > 200 and vals:
            break
    value = float(np.mean(vals))
    sample_txt = last["sample"] + f"; iteration count from {it_src}"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": len(vals), "warmup": args.warmup, "ms_per_step": 1e3 / value, "higher_is_better": True,
        "scaling": "strong" if int(os.environ.get("WORLD_SIZE", "1")) > 1 else "weak",
        "vs_baseline": None, "dtype": "f64 (PCG) / f32 (GNN)", "data": "synthetic",
        "config": {"workload": workload_text(g, N, E, args.rtol), "systems_per_step": 1},
        "parallelism": f"CPU: {last['cores']} torch threads for the GNN, 1 thread for scipy PCG (csr_matvec is single-threaded)",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": last["cores"], "kind": "port", "sample": sample_txt},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)



# ---------------------------------------------------------------------------------------------------------
# N > 1 (and --mode rowpart): ONE system row-partitioned over all ranks (BASELINE config C5; strong scaling)
# ---------------------------------------------------------------------------------------------------------
ROWPART_INPUTS = ("ei_r", "ei_t", "mv_r", "ea_t", "x_ext", "mask_ext")


def main_rowpart(args, ctx: Ctx, emit: bool = True):
    """The SAME 216^3 system at every N, split by rows: every rank holds only ITS slices of the inputs (the edges whose row
    it owns, the edges whose target it owns, the node window they touch -- parallel.rank_edge_slices) and computes only its
    share: the GNN on the target-owned edges (node-feature halo per layer), one boundary swap of G values, assembly of its
    rows of A, G and G^T, and the row-partitioned PCG (one persistent kernel per rank, csrc/pcg_fused.cuh: NVLink
    peer-store halos and the all-reduce of the CG scalars inside its grid barriers).  value = solves/s of that one system
    with the slices resident in HBM; e2e = the same with every rank's slices in pinned host memory (H2D of the slices and
    D2H of the owned part of x inside the timed region)."""
    from learningsparsepreconditioner4gpu_b200 import _lib, parallel as par, systems

    rank, world = ctx.rank, ctx.world
    barrier, max_over_ranks = ctx.barrier, ctx.max
    g = args.grid
    t0 = time.time()
    sample = systems.poisson_sample(g, seed=100, gen_device="cuda", device="cuda")   # generated once per rank, then sliced
    N, E = sample.num_nodes, sample.num_edges
    ws = systems.default_workspace(sample)
    b = ws.block_size
    starts = par.partition_rows(par.row_weights_from_edges(sample.edge_index, N), world)
    rs, re = starts[rank], starts[rank + 1]
    e0, e1, ids_t = par.rank_edge_slices(sample.edge_index, starts, rank)
    inp = {"ei_r": sample.edge_index[:, e0:e1].contiguous(), "ei_t": sample.edge_index[:, ids_t].contiguous(),
           "mv_r": sample.matrix_values[e0:e1].contiguous(), "ea_t": sample.edge_attr[ids_t].contiguous()}
    ps0 = par.PartitionedSystem(None, N, b, starts, rank=rank, world=world, slices=(inp["ei_r"], inp["ei_t"]))
    inp["x_ext"] = ps0.ext_slice(sample.x)
    inp["mask_ext"] = ps0.ext_slice(sample.mask.reshape(N, b))
    info0 = {"ext_nodes": ps0.ext_nodes, "edges_row_owned": int(e1 - e0), "edges_target_owned": int(ids_t.numel())}
    ps0.close()
    del sample, ps0, ids_t
    torch.cuda.empty_cache()
    host = {k: v.to("cpu").pin_memory() for k, v in inp.items()}
    h2d = tensor_bytes(*host.values())
    d2h = (re - rs) * b * 8
    log(f"[rowpart r{rank}] {g}^3: N={N} rows [{rs},{re}), slices {h2d / 1e6:.0f} MB, prepared in {time.time() - t0:.1f}s")

    def step(d, timers=None, profile=False):
        def mark(name):
            if timers is not None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                timers.append((name, ev))
        mark("start")
        ps = par.PartitionedSystem(None, N, b, starts, rank=rank, world=world, slices=(d["ei_r"], d["ei_t"]))
        mark("partition (layout, local edge lists, G-exchange plan)")
        g_t = ps.gnn_forward_local(ws.gnn, d["x_ext"], d["ea_t"])
        mark("gnn (target-owned edges, halo per layer)")
        local = ps.assemble_local(d["mv_r"], g_t, d["mask_ext"])
        rhs_loc = local[0].matvec(d["mask_ext"].reshape(-1).to(torch.float64))
        mark("G swap + assemble A, G, G^T rows")
        solver = ps.make_solver(local, ws.epsilon, reuse=True)   # the IPC connection is kept between steps
        mark("connect / relayout")
        x, res, _ = solver.solve(rhs_loc, "ext_spai", rtol=args.rtol)
        mark("pcg (row-partitioned)")
        info = solver.info()
        info["pcg_mode"] = int(res.reserved)
        if profile and res.reserved == 4:
            ms, its = solver.fused_profile()
            info["fused_phases_ms_per_iteration"] = {k: round(v / max(its, 1), 5) for k, v in zip(FUSED_PHASES, ms)}
            if world > 1:
                info["cross_rank_barrier"] = solver.fused_barrier_profile()
        if profile:
            info["nnz_local"] = [local[0].nnz, local[1].nnz, local[2].nnz]
            info["stream_bytes"] = [list(local[0].stream_bytes()), list(local[1].stream_bytes())]
        solver.close()      # parks the connection for the next step
        ps.close()
        return x, res, info

    for _ in range(max(args.warmup, 3)):
        x, res, info = step(inp, profile=True)
    launches0 = _lib.total_launches()
    sampler = ClockSampler(ctx.local)
    timers = []
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    pcg_ms, iters = [], []
    for _ in range(args.steps):
        x, res, _ = step(inp, timers)
        pcg_ms.append(float(res.solve_ms))
        iters.append(int(res.iterations))
    ev1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    pcg = max_over_ranks(float(np.mean(pcg_ms)))
    ph = phases_ms(timers)
    launches = _lib.total_launches() - launches0

    # ---- end to end: every rank's slices start in pinned host memory ----
    x_host = torch.empty(max((re - rs) * b, 1), dtype=torch.float64).pin_memory()

    def e2e_step():
        d = {k: v.to("cuda", non_blocking=True) for k, v in host.items()}
        xs, r, _ = step(d)
        x_host[: xs.numel()].copy_(xs, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return r

    e2e_step()
    barrier()
    ev0.record()
    for _ in range(args.steps):
        e2e_step()
    ev1.record()
    barrier()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1))
    h2d_all, d2h_all = ctx.sum(float(h2d)), ctx.sum(float(d2h))

    # ---- every rank's view of the iteration (who waits for whom): phases, barrier waits, local sizes ----
    mine = {"rank": rank, "rows": int(re - rs), "nnz_local": info.get("nnz_local"), "phases_ms_per_iteration": info.get("fused_phases_ms_per_iteration"),
            "cross_rank_barrier": info.get("cross_rank_barrier")}
    per_rank = [mine]
    if world > 1:
        import torch.distributed as dist
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)
    # ---- roofline of the dominant phase: the three matrix sweeps of this rank's share ----
    peak, peak_src = peaks()
    n_loc = (re - rs) * b
    roofline = None
    fp = info.get("fused_phases_ms_per_iteration")
    if fp:
        nA, nG, nGT = info["nnz_local"]
        (va, ia), (vg, ig) = info["stream_bytes"]
        ra_, rg_ = (8 if ia == 2 else 4), (8 if ig == 2 else 4)
        streamed = [(vg + ig) * nGT + (rg_ + 16) * n_loc, (vg + ig) * nG + (rg_ + 24) * n_loc, (va + ia) * nA + (ra_ + 16) * n_loc]
        alg = [12 * nGT + 20 * n_loc + 4, 12 * nG + 28 * n_loc + 4, 12 * nA + 20 * n_loc + 4]
        sw = [fp["GT_r+barrier"], fp["G_t+barrier"], fp["A_p+barrier"]]
        ach = sum(streamed) / (sum(sw) * 1e-3) / 1e9
        roofline = {"kernel": "pcg_fused_kernel<4096,2,8,float,u16,4>: the three matrix sweeps of one iteration on rank 0's rows, each timed "
                              "INCLUDING the cross-rank barrier (halo exchange + all-reduce) that closes it",
                    "bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4), "peak_source": peak_src,
                    "traffic": None, "bytes_per_sweep_rank0": [int(v) for v in streamed], "ms_per_sweep_rank0": sw,
                    "algorithmic": {"bytes_per_sweep_rank0": alg, "GB/s": round(sum(alg) / (sum(sw) * 1e-3) / 1e9, 1)},
                    "phases_ms_per_iteration_rank0": fp, "cross_rank_barrier_rank0": info.get("cross_rank_barrier")}
    line = None
    if rank == 0:
        W = workload_text(g, N, E, args.rtol)
        line = {
            "metric": METRIC, "value": round(args.steps / (dev_ms * 1e-3), 4), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(dev_ms / args.steps, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64 (PCG) / f32 (GNN)", "data": "synthetic",
            "config": {"workload": W, "systems_per_step": 1},
            "parallelism": f"ONE system row-partitioned over {world} GPU(s): GNN on target-owned edges (NCCL node-feature halo per layer), "
                           f"owned rows of A / G / G^T assembled per rank, PCG in one persistent kernel per rank with NVLink peer-store "
                           f"halos and peer-memory all-reduce of the CG scalars",
            "l2": "per-rank matrices larger than L2 up to 8 ranks (A + G + G^T = %.2f GB per rank)" % ((E * 3 * 6 + 24.0 * N) / world / 1e9),
            "time_to_tol_ms": round(pcg, 3), "iterations": float(np.mean(iters)), "converged": int(res.converged),
            "ms_per_iteration": round(pcg / max(float(np.mean(iters)), 1.0), 4), "pcg_mode": info.get("pcg_mode"),
            "phases_ms_per_step": {k: round(v / args.steps, 3) for k, v in ph.items()},
            "halo": {k: info[k] for k in ("n_push", "n_recv", "push_elems", "owned_offset") if k in info}, "rank0": info0,
            "e2e": {"value": round(args.steps / (e2e_ms * 1e-3), 4), "unit": UNIT, "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                    "ms_per_step": round(e2e_ms / args.steps, 3),
                    "api": "parallel.PartitionedSystem / DistPCG on every rank's pinned-host slices of the system (bytes summed over ranks)"},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": None, "clocks": clocks, "per_rank": per_rank,
        }
        if emit:
            print(json.dumps(line), flush=True)
    barrier()
    par.DistPCG.drop_connections()
    return line


# ---------------------------------------------------------------------------------------------------------
# --mode batch_small: BASELINE config C2 -- a batch of small independent Poisson systems per GPU
# ---------------------------------------------------------------------------------------------------------
def main_batch_small(args):
    """Every rank solves its own shard of a batch of independent systems (N between 4 k and 32 k nodes, the size range
    of the reference's poisson3d_tetmesh set, preprocess/msh_to_npy.py:77-86), one after the other through
    ``workspace.solve``: GNN -> G -> assembly -> PCG to rtol.  No data-path collective; solves/s over all ranks."""
    from learningsparsepreconditioner4gpu_b200 import _lib, systems

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    _lib.require_cuda()
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    grids = [16, 20, 25, 32]                         # 4096, 8000, 15625, 32768 nodes
    per_rank = max(args.batch // world, 1)
    samples = [systems.poisson_sample(grids[(rank * per_rank + k) % 4], seed=1000 + rank * per_rank + k, device="cuda")
               for k in range(per_rank)]
    K = max(int(args.streams), 1)
    # K systems in flight per GPU: one workspace + one non-blocking CUDA stream per host thread (batch.BatchSolver)
    from learningsparsepreconditioner4gpu_b200.batch import BatchSolver

    solver = BatchSolver(lambda: systems.default_workspace(samples[0]), streams=K)
    # sequential reference pass: the concurrent passes must reproduce its iteration counts
    ref_iters = []
    for s_ in samples:
        x, res, _, A, G = solver.workspaces[0].solve(s_, rtol=args.rtol)
        ref_iters.append(int(res.iterations))
        del A, G
    torch.cuda.synchronize()

    collated = None
    if args.collated:
        # the batch as ONE block-diagonal system: one GNN forward / assembly / G^T for all systems, then every block solved
        # by its own CTA group inside one launch per wave (batch.CollatedBatchSolver -> b200spai_pcg_solve_batch)
        from learningsparsepreconditioner4gpu_b200.batch import CollatedBatchSolver

        collated = CollatedBatchSolver(solver.workspaces[0])

    def one_pass():
        if collated is not None:
            results = collated.solve(samples, rtol=args.rtol)
            its = [int(r.iterations) for _, r in results]
            # same arithmetic per system; the partial sums of its dot products are grouped by another CTA count
            assert all(r.converged == 1 for _, r in results), its
            assert all(abs(a - b) <= 1 + b // 50 for a, b in zip(its, ref_iters)), (its, ref_iters)
            return sum(its)
        results = solver.solve(samples, rtol=args.rtol)
        its = [int(r.iterations) for _, r in results]
        assert all(r.converged == 1 for _, r in results) and its == ref_iters, (its, ref_iters)
        return sum(its)

    for _ in range(max(args.warmup, 1)):
        one_pass()
    launches0 = _lib.total_launches()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    its = 0
    for _ in range(args.steps):
        its += one_pass()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    if rank == 0:
        total = per_rank * world * args.steps
        print(json.dumps({
            "metric": "learned-SPAI PCG solves/s to rtol 1e-6, batch of small independent systems (BASELINE config C2)",
            "mode": "batch_small", "value": round(total / (ms * 1e-3), 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 1), "ms_per_step": round(ms / args.steps, 3), "ms_per_system": round(ms / args.steps / per_rank, 3),
            "higher_is_better": True, "scaling": "weak", "dtype": "f64 (PCG) / f32 (GNN)", "data": "synthetic",
            "config": {"workload": f"{per_rank} Poisson systems per GPU, N in {{4096, 8000, 15625, 32768}}, Dirichlet patch, "
                                   f"random-init GNN, rhs = A*mask, rtol {args.rtol:g}; " +
                                   ("the batch collated into one block-diagonal system (one GNN forward / assembly per batch, "
                                    "one PCG launch per wave of systems, a CTA group per system)" if args.collated else
                                    f"{K} system(s) in flight per GPU (host threads x CUDA streams)")},
            "collated": bool(args.collated),
            "mean_iterations": round(its / (per_rank * args.steps), 1),
            "gpu_launches": int(_lib.total_launches() - launches0)}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------------
# --mode shapes: BASELINE configs C1 (heat), C3 (elasticity, b = 3), C4 (synthetic SPD, irregular rows) at their sizes
# ---------------------------------------------------------------------------------------------------------
def main_shapes(args):
    """One JSON line per named shape (systems.SHAPES): the whole path (graph preprocessing, GNN, assembly, G^T, PCG to
    rtol) per step on ONE GPU, per-kernel GB/s of the PCG iteration, row-length histogram, and beside it the REAL oracle
    run of the same system on this box's host cores (oracle GNN forward + the scipy-cg recurrence to convergence): its
    time is the cpu_baseline, its iteration count and G are the parity columns."""
    from learningsparsepreconditioner4gpu_b200 import _lib, systems
    from learningsparsepreconditioner4gpu_b200.device import DevicePCG

    _lib.require_cuda()
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:      # N GPUs: one independent system of the shape per GPU (weak scaling, no data-path collective)
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peak, peak_src = peaks()
    names = args.shapes.split(",") if args.shapes else list(systems.SHAPES)
    for name in names:
        cpu_s, b = systems.shape_sample(name)
        sample = cpu_s.to("cuda")
        ws = systems.default_workspace(cpu_s, block_size=b)
        N, E, n = cpu_s.num_nodes, cpu_s.num_edges, cpu_s.num_nodes * b
        for _ in range(max(args.warmup, 3)):
            x, res, solver, A, G, rhs = run_step(ws, sample, args.rtol)
            del solver, A, G, rhs
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()
        launches0 = _lib.total_launches()
        timers = []
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        pcg_ms, iters = [], []
        for k in range(args.steps):
            x, res, solver, A, G, rhs = run_step(ws, sample, args.rtol, timers)
            pcg_ms.append(float(res.solve_ms))
            iters.append(int(res.iterations))
            if k < args.steps - 1:
                del solver, A, G, rhs
        ev1.record()
        torch.cuda.synchronize()
        dev_ms = ev0.elapsed_time(ev1)
        if dist is not None:     # max over ranks
            tmax = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dev_ms = float(tmax.item())
        launches = _lib.total_launches() - launches0
        ph = phases_ms(timers)
        mode_ran = int(res.reserved)
        true_rel = float((rhs - A.matvec(x)).norm() / rhs.norm())
        # per-kernel times of the 5-launch iteration on these matrices (the kernels mode 3 / 4 fuse)
        kms = solver.profile_iterations(rhs, "ext_spai", iterations=50)
        nnzA, nnzG = A.nnz, G.nnz
        (va, ia), (vg, ig) = A.stream_bytes(), G.stream_bytes()
        alg = [12 * nnzG + 20 * n + 4, 12 * nnzG + 28 * n + 4, 24 * n, 12 * nnzA + 20 * n + 4, 48 * n]
        knames = ["GT_r", "G_t", "update_p", "A_p", "update_xr"]
        per_kernel = {k: {"ms": round(ms, 5), "algorithmic_bytes": by, "GB/s": round(by / (ms * 1e-3) / 1e9, 1) if ms > 0 else None}
                      for k, by, ms in zip(knames, alg, kms)}
        spmv_ms = (kms[0] + kms[1] + kms[3]) / 3
        spmv_bytes = (alg[0] + alg[1] + alg[3]) / 3
        rl = np.diff(A.to_scipy().indptr)
        hist_edges = [1, 2, 4, 8, 16, 32, 64, 128, 1 << 30]
        row_hist = {f"<{hi}": int(((rl >= lo) & (rl < hi)).sum()) for lo, hi in zip([0] + hist_edges[:-1], hist_edges)}
        # ---- the oracle beside it: a REAL run of the same system on the host cores ----
        cpu = None
        parity = None
        if not args.no_cpu_baseline and rank == 0 and world == 1:
            from oracle import assemble as oasm, gnn as ognn, pcg as opcg

            sd = {k: v.detach().cpu() for k, v in ws.state_dict().items()}
            t0 = time.perf_counter()
            Go = ognn.workspace_forward(sd, cpu_s.x, cpu_s.edge_index, cpu_s.edge_attr, b)
            t_gnn = time.perf_counter() - t0
            Ao = oasm.to_csr_cpu(cpu_s.edge_index, cpu_s.matrix_values, n, cpu_s.mask)
            Gc = oasm.to_csr_cpu(cpu_s.edge_index, Go, n, cpu_s.mask)
            gt = cpu_s.mask.numpy().reshape(-1).astype(np.float64)
            timer = opcg.ScipyPcgTimer(Ao, Gc, ws.epsilon)
            its_o, t_pcg = timer.run(gt, args.rtol)
            cpu = {"value": round(1.0 / (t_gnn + t_pcg), 5), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                   "sample": f"the whole system: oracle GNN forward {t_gnn:.2f} s ({torch.get_num_threads()} torch threads) + oracle scipy "
                             f"PCG {its_o} iterations in {t_pcg:.2f} s (1 thread); not extrapolated",
                   "gnn_s": round(t_gnn, 3), "pcg_s": round(t_pcg, 3), "iterations": int(its_o)}
            with torch.no_grad():
                Gd = ws(sample.x, sample.edge_index, sample.edge_attr).cpu().numpy()
            As = A.to_scipy()
            parity = {"oracle_iterations": int(its_o), "cuda_iterations": int(np.mean(iters)),
                      "dG_over_maxG": float(np.abs(Gd - Go.numpy()).max() / np.abs(Go.numpy()).max()),
                      "pattern_A_bit_exact": bool(np.array_equal(As.indptr, Ao.indptr) and np.array_equal(As.indices, Ao.indices)
                                                  and np.array_equal(As.data, Ao.data))}
        line = {
            "metric": METRIC, "mode": "shapes", "shape": name, "value": round(world * args.steps / (dev_ms * 1e-3), 3), "unit": UNIT,
            "n_gpus": world, "parallelism": f"{world} independent system(s) of this shape, one per GPU", "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(dev_ms / args.steps, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 (PCG) / f32 (GNN)", "data": "synthetic",
            "config": {"workload": f"{name}: N = {N} nodes, block {b}, n = {n} rows, E = {E} stored blocks, nnz(A) = {nnzA}, "
                                   f"nnz(G) = {nnzG}; random-init GNN (seed 42), eps 3e-3, rhs = A*mask, rtol {args.rtol:g}",
                       "l2": "whole system is L2-resident (%.1f MB): latency-bound, GB/s are L2 rates" % ((nnzA + 2 * nnzG) * 12 / 1e6)},
            "time_to_tol_ms": round(float(np.mean(pcg_ms)), 3), "iterations": float(np.mean(iters)), "converged": int(res.converged),
            "true_rel_residual": true_rel, "pcg_mode": mode_ran,
            "ms_per_iteration": round(float(np.mean(pcg_ms)) / max(float(np.mean(iters)), 1.0), 5),
            "phases_ms_per_step": {k: round(v / args.steps, 3) for k, v in ph.items()},
            "row_length": {"min": int(rl.min()), "mean": round(float(rl.mean()), 2), "max": int(rl.max()), "histogram": row_hist},
            "stream_bytes_per_entry": {"A": [va, ia], "G": [vg, ig]},
            "roofline": {"kernel": "SpMV family of the 5-launch iteration (profile_iterations), mean of t = G^T r, z = G t + eps r, q = A p",
                         "bound": "hbm", "achieved": round(spmv_bytes / (spmv_ms * 1e-3) / 1e9, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(spmv_bytes / (spmv_ms * 1e-3) / 1e9 / peak, 4), "peak_source": peak_src, "traffic": None,
                         "per_kernel": per_kernel},
            "gpu_launches": int(launches), "cpu_baseline": cpu, "parity": parity,
        }
        if rank == 0:
            print(json.dumps(line), flush=True)
        del solver, A, G, rhs
        torch.cuda.empty_cache()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------------
# --mode train: one optimiser step through the fused kernels (the reference's training_step)
# ---------------------------------------------------------------------------------------------------------
def main_train(args):
    """A step = training_step (GNN forward with tape -> AATPE -> RelativeL2Loss_ANorm) + backward + clip + AdamW on ONE
    collated batch of `--batch` Poisson systems (reference defaults: batch_size 4, config/basic.yaml:9; adamw lr 1e-3,
    gradient_clip_val 10, config/trainer.yaml; meshes of 4 k - 32 k nodes like its poisson3d_tetmesh set).  Under torchrun:
    data parallel, every rank its own batches, one all-reduce of the 70 KB gradient per step (weak scaling)."""
    from learningsparsepreconditioner4gpu_b200 import _lib, systems
    from learningsparsepreconditioner4gpu_b200.neural_cg.data import collate

    from learningsparsepreconditioner4gpu_b200.neural_cg.trainer import allreduce_mean_grads

    _lib.require_cuda()
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:   # data parallel: every rank its own batches, ONE all-reduce of the flat gradient per step (inside the timed step)
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = max(min(args.batch, 16), 1) if args.batch != 64 else 4
    grids = [16, 20, 25, 32]
    gen = torch.Generator().manual_seed(rank)
    batches_host = []
    for k in range(4):                                   # four different batches, cycled: no step sees a cached input
        samples = [systems.poisson_sample(grids[(k + j) % 4], seed=2000 + 100 * rank + 10 * k + j, device="cpu") for j in range(B)]
        for s_ in samples:
            s_.residual = torch.randn(s_.num_nodes, 1, generator=gen) * s_.mask
        batches_host.append(collate(samples).pin_memory())
    dev_batches = [b.to("cuda") for b in batches_host]
    ws = systems.default_workspace(dev_batches[0]).cuda().train()
    opt = torch.optim.AdamW(ws.parameters(), lr=1e-3, weight_decay=3e-3)
    params = [p for p in ws.parameters() if p.requires_grad]

    def step(batch):
        opt.zero_grad(set_to_none=True)
        loss = ws.training_step(batch, 0)
        loss.backward()
        allreduce_mean_grads(params)                     # no-op on one GPU
        torch.nn.utils.clip_grad_norm_(params, 10.0)
        opt.step()
        return loss.detach()

    for k in range(max(args.warmup, 3)):
        step(dev_batches[k % 4])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > the 126 MB L2, rewritten between steps
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = _lib.total_launches()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    losses = []
    for k in range(args.steps):
        flush.fill_(k & 255)
        ev[k][0].record()
        losses.append(step(dev_batches[k % 4]))
        ev[k][1].record()
    torch.cuda.synchronize()
    launches = _lib.total_launches() - launches0
    ms = sum(a.elapsed_time(b) for a, b in ev)
    # end to end: the batch comes from pinned host memory every step, the loss goes back to the host
    h2d = sum(tensor_bytes(*[v for v in batches_host[k % 4].__dict__.values() if torch.is_tensor(v)]) for k in range(args.steps))
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0.record()
    for k in range(args.steps):
        flush.fill_(k & 255)
        loss_host = float(step(batches_host[k % 4].to("cuda", non_blocking=True)).cpu())
    t1.record()
    torch.cuda.synchronize()
    e2e_ms = t0.elapsed_time(t1)
    clocks = sampler.stop()
    if dist is not None:   # the slowest rank sets the pace
        tmax = torch.tensor([ms, e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        ms, e2e_ms = float(tmax[0]), float(tmax[1])
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return
    nodes = sum(int(b.num_nodes) for b in batches_host) / 4.0
    edges = sum(int(b.num_edges) for b in batches_host) / 4.0
    # algorithmic flops of the GNN per step: forward 2 * (23.6 k per edge + 7.3 k per node) MAC-flops,
    # backward = recompute + input gradients + weight gradients = 3 x forward
    gnn_flops = 4.0 * (edges * 23.6e3 + nodes * 7.3e3)
    print(json.dumps({
        "metric": "training samples/s through the fused kernels (training_step + backward + clip + AdamW; SURVEY 8f-4)",
        "mode": "train", "value": round(world * B * args.steps / (ms * 1e-3), 2), "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"collated batches of {B} jittered Kuhn-tet Poisson systems, N in {{4096, 8000, 15625, 32768}} "
                               f"(mean {nodes:.0f} nodes, {edges:.0f} edges per batch), Dirichlet patch, random rhs, "
                               "RelativeL2Loss_ANorm, AdamW lr 1e-3, clip 10; L2 flushed between steps (256 MB fill)"},
        "parallelism": f"dp{world}: every rank its own batches, one all-reduce of the flat gradient (70 KB) per step" if world > 1 else "1 GPU",
        "e2e": {"value": round(world * B * args.steps / (e2e_ms * 1e-3), 2), "unit": "samples/s",
                "h2d_bytes_per_step": int(h2d / args.steps), "d2h_bytes_per_step": 4},
        "roofline": {"bound": "fp32 FFMA", "achieved": round(gnn_flops / (ms / args.steps * 1e-3) / 1e12, 2), "peak": 74.5,
                     "unit": "TFLOP/s", "frac": round(gnn_flops / (ms / args.steps * 1e-3) / 74.5e12, 3), "traffic": None,
                     "note": "per GPU; algorithmic GNN flops of forward + backward (4 x the forward MAC-flops) over the WHOLE step incl. "
                             "loss, optimiser and host launch gaps; peak = 148 SMs x 128 FFMA x 1.965 GHz"},
        "loss_first_last": [float(losses[0]), float(losses[-1])], "loss_e2e_last": loss_host,
        "gpu_launches": int(launches), "clocks": clocks}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--grid", type=int, default=216)
    ap.add_argument("--rtol", type=float, default=1e-6)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--shapes", default="", help="shapes: comma-separated subset of heat,elast,synth10240,synth15360,synth20480")
    ap.add_argument("--no-replicas", action="store_true", help="N > 1: skip the extra independent-systems (replicas) figure")
    ap.add_argument("--mode", default="auto", choices=["auto", "batch", "rowpart", "batch_small", "shapes", "train"],
                    help="auto (the bench contract): N = 1 -> one 10 M-DOF system on one GPU; N > 1 -> the SAME system "
                         "row-partitioned over the N GPUs (strong scaling) with the independent-replicas figure as an extra key; "
                         "batch: one independent system per GPU (weak scaling); rowpart: the row-partitioned line at any N; "
                         "batch_small: a batch of small systems per GPU (BASELINE config C2); shapes: C1 / C3 / C4 at their sizes")
    ap.add_argument("--batch", type=int, default=64, help="batch_small: systems per step over all ranks")
    ap.add_argument("--streams", type=int, default=4, help="batch_small: systems in flight per GPU (host threads x CUDA streams)")
    ap.add_argument("--no-collated", dest="collated", action="store_false",
                    help="batch_small: host threads x CUDA streams, one system per workspace.solve (default: the batch collated into "
                         "one block-diagonal system, one PCG launch per wave of systems)")
    ap.set_defaults(collated=True)
    args = ap.parse_args()
    if args.impl == "reference":
        main_reference(args)
    elif args.mode == "batch_small":
        main_batch_small(args)
    elif args.mode == "shapes":
        main_shapes(args)
    elif args.mode == "train":
        main_train(args)
    else:
        ctx = Ctx()
        if args.mode == "rowpart" or (args.mode == "auto" and ctx.world > 1):
            line = main_rowpart(args, ctx, emit=False)
            if ctx.world > 1 and not args.no_replicas and args.mode == "auto":
                # the round-1 headline, kept beside the strong-scaling line: N independent systems, one per GPU
                import copy

                a2 = copy.copy(args)
                a2.steps, a2.no_cpu_baseline = min(args.steps, 2), True
                rep = main_ours(a2, ctx, emit=False)
                if line is not None and rep is not None:
                    line["replicas"] = {k: rep[k] for k in ("value", "unit", "ms_per_step", "steps", "iterations", "ms_per_iteration",
                                                            "parallelism")}
                    line["replicas"]["scaling"] = "weak"
                    line["replicas"]["e2e"] = rep["e2e"]
            if line is not None:
                print(json.dumps(line), flush=True)
        else:
            main_ours(args, ctx)
        ctx.close()


if __name__ == "__main__":
    main()
