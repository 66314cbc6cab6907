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
# PCG iterations of the default workload at rtol 1e-6, measured by the CUDA path (iteration parity with the oracle
# is +-1 on every size the oracle can solve, tests/test_pcg_gpu.py).  Used ONLY by --impl reference to scale its
# per-iteration CPU time to a whole solve without running ~10^3 s of scipy; override with B200SPAI_REF_ITERS.
DEFAULT_ITERS = {216: 1391}


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


def main_ours(args):
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
    n_gpus = world
    if args.gpus and args.gpus != world and rank == 0:
        log(f"[bench] --gpus {args.gpus} but WORLD_SIZE={world}: running on {world} rank(s)")

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v: float) -> float:
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

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
    achieved = spmv_bytes / 3 / (spmv_ms / 3 * 1e-3) / 1e9
    streamed = spmv_streamed / 3 / (spmv_ms / 3 * 1e-3) / 1e9
    iter_bytes = sum(bytes_k.values())
    iter_streamed = sum(streamed_k.values())
    iter_ms = sum(kms)
    roofline = {
        "kernel": "spmv_stream_kernel<4096,2,8,float,u16,*> (3 launches per PCG iteration: t=G^T r, z=G t+eps r, q=A p)",
        "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
        "peak_source": peak_src,
        "traffic": 1.11e9 if g == 216 else None,
        "traffic_source": "ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch at 216^3 (profiles/r01_spmv_stream_ncu_full.txt)",
        "bytes_per_launch": spmv_bytes // 3, "ms_per_launch": round(spmv_ms / 3, 4),
        "note": ("achieved = ALGORITHMIC bytes (SURVEY 8d: 12 B per entry, fp64 value + i32 index) / time; the kernel streams a lossless "
                 f"{per_a} B per entry (fp32 values widened on load, 16-bit column steps), so it can exceed the copy peak; "
                 "'streamed' is the same time against the bytes really read"),
        "streamed": {"bytes_per_launch": spmv_streamed // 3, "GB/s": round(streamed, 1), "frac": round(streamed / peak, 4)},
        "pcg_iteration": {"ms": round(iter_ms, 4), "bytes": iter_bytes, "GB/s": round(iter_bytes / (iter_ms * 1e-3) / 1e9, 1),
                          "frac": round(iter_bytes / (iter_ms * 1e-3) / 1e9 / peak, 4),
                          "streamed_bytes": iter_streamed, "streamed_GB/s": round(iter_streamed / (iter_ms * 1e-3) / 1e9, 1),
                          "streamed_frac": round(iter_streamed / (iter_ms * 1e-3) / 1e9 / peak, 4)},
        "per_kernel": detail,
    }
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
            "config": {"workload": f"poisson3d jittered Kuhn-tet grid {g}^3 = {N} DOF per GPU, E = {E} stored entries, "
                                   f"Dirichlet patch, random-init GNN (seed 42), eps 3e-3, rhs = A*mask, rtol {args.rtol:g}",
                       "systems_per_step": n_gpus, "parallelism": f"{n_gpus} independent system(s), one per GPU",
                       "l2": "inputs larger than L2 (A + G + G^T = %.1f GB)" % ((nnzA + 2 * nnzG) * 12 / 1e9)},
            "time_to_tol_ms": round(float(np.mean(pcg_ms)), 3), "iterations": round(total_iters, 1), "converged": converged,
            "ms_per_iteration": round(float(np.mean(pcg_ms)) / max(float(np.mean(iters)), 1.0), 4),
            "pcg_mode": pcg_mode, "fused_phases": fused_phases,
            "phases_ms_per_step": {k: round(v / args.steps, 3) for k, v in ph.items()},
            "e2e": {"value": round(e2e_value, 4), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": round(e2e_ms / args.steps, 3), "api": "SimpleTrainingWorkspace.solve path on a pinned-host sample"},
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


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
    arm = CpuArm(sample.to("cpu"), ws, args.rtol, gnn_grid=56)
    arm.step(iterations, 1)  # warm-up (page in, torch thread pool)
    return arm.step(iterations, 6)


def main_reference(args):
    """--impl reference: the reference's CPU path (oracle port; the reference itself needs torch_geometric,
    lightning and pymathprim, none installable here) on this box's host cores, same workload and metric."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from learningsparsepreconditioner4gpu_b200 import systems

    g = args.grid
    iterations = int(os.environ.get("B200SPAI_REF_ITERS", DEFAULT_ITERS.get(g, 0)) or 0)
    if iterations <= 0:
        iterations = int(6.4 * g)  # measured: 1372 PCG iterations at g = 216 (random-init G); linear in the grid edge
    gen_device = "cuda" if torch.cuda.is_available() else "cpu"   # input generation only
    sample = systems.poisson_sample(g, seed=100, gen_device=gen_device, device="cpu")
    ws = systems.default_workspace(sample)
    N, E = sample.num_nodes, sample.num_edges
    arm = CpuArm(sample, ws, args.rtol, gnn_grid=40)
    steps = max(args.steps, 1)
    vals, last = [], None
    t_begin = time.perf_counter()
    for i in range(max(args.warmup, 0) + steps):
        # every step is a bounded sample: a few PCG iterations on the full system + a small-grid GNN forward
        last = arm.step(iterations, pcg_iters=3)
        if i >= args.warmup:
            vals.append(last["value"])
        if time.perf_counter() - t_begin > 200 and vals:
            break
    value = float(np.mean(vals))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", "1")),
        "steps": len(vals), "warmup": args.warmup, "ms_per_step": 1e3 / value, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64 (PCG) / f32 (GNN)", "data": "synthetic",
        "config": {"workload": f"poisson3d jittered Kuhn-tet grid {g}^3 = {N} DOF, E = {E} stored entries, Dirichlet patch, "
                               f"random-init GNN (seed 42), eps 3e-3, rhs = A*mask, rtol {args.rtol:g}",
                   "systems_per_step": 1, "parallelism": "CPU: torch threads for the GNN, 1 thread for scipy PCG"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": last["cores"], "kind": "port", "sample": last["sample"]},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)



# ---------------------------------------------------------------------------------------------------------
# --mode rowpart: ONE system row-partitioned over all ranks (BASELINE config C5; strong scaling)
# ---------------------------------------------------------------------------------------------------------
def main_rowpart(args):
    """Every rank holds the INPUTS of the same system and computes only its share (parallel.PartitionedSystem): the GNN
    on the edges whose target it owns (node-feature halo per layer), one boundary swap of G values, assembly of its rows
    of A, G and G^T, and the row-partitioned PCG (csrc/dist.cu: NVLink peer-store halos, peer-memory all-reduce)."""
    from learningsparsepreconditioner4gpu_b200 import _lib, parallel as par, systems
    from learningsparsepreconditioner4gpu_b200.neural_cg.nn.gnns import DeviceGraph
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    _lib.require_cuda()
    torch.cuda.set_device(local)
    import torch.distributed as dist

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local), rank=rank, world_size=world)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    g = args.grid
    sample = systems.poisson_sample(g, seed=100, gen_device="cuda", device="cuda")   # the SAME inputs on every rank
    N, E = sample.num_nodes, sample.num_edges
    ws = systems.default_workspace(sample)
    starts = par.partition_rows(par.row_weights_from_edges(sample.edge_index, N), world)
    rs, re = starts[rank], starts[rank + 1]
    log(f"[rowpart r{rank}] {g}^3: N={N} rows [{rs},{re})")

    def step(timers=None):
        def mark(name):
            if timers is not None:
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                timers.append((name, ev))
        mark("start")
        ps = par.PartitionedSystem(sample.edge_index, N, ws.block_size, starts)
        mark("partition (edge ownership, local edge lists)")
        g_t = ps.gnn_forward(ws.gnn, sample.x, sample.edge_attr)
        mark("gnn (target-owned edges, halo per layer)")
        local = ps.assemble(sample.matrix_values, g_t, sample.mask)
        rhs_loc = local[0].matvec(ps.ext_slice(sample.mask.reshape(N, ws.block_size)).reshape(-1).to(torch.float64))
        mark("G swap + assemble A, G, G^T rows")
        solver = ps.make_solver(local, ws.epsilon, reuse=True)   # the IPC connection is kept between steps
        mark("connect / relayout")
        x, res, _ = solver.solve(rhs_loc, "ext_spai", rtol=args.rtol)
        mark("pcg (row-partitioned)")
        info = solver.info()
        if timers is None and os.environ.get("B200SPAI_DIST_PROFILE"):     # warm-up steps only: per-launch times
            info["per_launch_ms"] = solver.profile_iterations(rhs_loc, 50)
        info["edges_gnn"] = int(ps.ids_t.numel())
        solver.close()      # parks the connection for the next step
        ps.close()
        return res, info

    for _ in range(max(args.warmup, 1)):
        res, info = step()
    prof = info.get("per_launch_ms")
    if prof:
        log(f"[rowpart r{rank}] per-launch ms: {prof}")
    launches0 = _lib.total_launches()
    timers = []
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    pcg_ms, iters = [], []
    for _ in range(args.steps):
        res, info = step(timers)
        pcg_ms.append(float(res.solve_ms))
        iters.append(int(res.iterations))
    ev1.record()
    barrier()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    pcg = max_over_ranks(float(np.mean(pcg_ms)))
    ph = phases_ms(timers)
    launches = _lib.total_launches() - launches0
    if rank == 0:
        line = {
            "metric": "row-partitioned learned-SPAI PCG, one system over all GPUs: solves/s (strong scaling)", "mode": "rowpart",
            "value": round(args.steps / (dev_ms * 1e-3), 4), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 1), "ms_per_step": round(dev_ms / args.steps, 3), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64 (PCG) / f32 (GNN)", "data": "synthetic",
            "config": {"workload": f"ONE poisson3d jittered Kuhn-tet system {g}^3 = {N} DOF, E = {E}, row-partitioned over {world} GPU(s); "
                                   f"GNN on target-owned edges, assembly of owned rows, PCG partitioned; rtol {args.rtol:g}",
                       "halo": info, "rows_rank0": [rs, re]},
            "time_to_tol_ms": round(pcg, 3), "iterations": float(np.mean(iters)), "converged": int(res.converged),
            "ms_per_iteration": round(pcg / max(float(np.mean(iters)), 1.0), 4),
            "phases_ms_per_step": {k: round(v / args.steps, 3) for k, v in ph.items()}, "gpu_launches": int(launches),
        }
        print(json.dumps(line), flush=True)
    barrier()
    par.DistPCG.drop_connections()
    dist.destroy_process_group()


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

    def one_pass():
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
                                   f"random-init GNN, rhs = A*mask, rtol {args.rtol:g}; {K} system(s) in flight per GPU "
                                   f"(host threads x CUDA streams)"},
            "mean_iterations": round(its / (per_rank * args.steps), 1),
            "gpu_launches": int(_lib.total_launches() - launches0)}), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--grid", type=int, default=216)
    ap.add_argument("--rtol", type=float, default=1e-6)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", default="batch", choices=["batch", "rowpart", "batch_small"],
                    help="batch: one independent 10 M-DOF system per GPU (the bench contract); rowpart: one system over all "
                         "GPUs; batch_small: a batch of small systems per GPU (BASELINE config C2)")
    ap.add_argument("--batch", type=int, default=64, help="batch_small: systems per step over all ranks")
    ap.add_argument("--streams", type=int, default=4, help="batch_small: systems in flight per GPU (host threads x CUDA streams)")
    args = ap.parse_args()
    if args.impl == "reference":
        main_reference(args)
    elif args.mode == "rowpart":
        main_rowpart(args)
    elif args.mode == "batch_small":
        main_batch_small(args)
    else:
        main_ours(args)


if __name__ == "__main__":
    main()
