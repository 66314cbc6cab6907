"""The reference's inference benchmark driver on the B200 path (reference ``infer.py``).

What is kept from ``infer.py``: ``InferenceTimestat`` / ``Timestat`` with the same keys, the same summary and
per-sample CSV schemas and file names (``infer_{prefix}{exp}_{-log10 rtol}.csv``, ``all_infer_...csv``, infer.py:372-384,
so ``misc/tab_to_latex_benchmark.py`` reads them unchanged), ``get_workspace`` (infer.py:153-163), the right-hand-side
modes ``mask`` / ``ones`` / ``random`` / ``neighbour`` (:297-307), ``repeat``, the GNN warm-up (:270-275), the per-sample
flow (A from ``to_csr_cpu``, preconditioner time = mean ``inference_step`` time, then the classical baselines and the
neural solve, :277-331).

What differs, on purpose: the solves run on ``cuda`` only -- keys ``PCG-none-cuda``, ``PCG-diagonal-cuda``,
``Neural+CUDA`` (the reference's ``*-cpu`` rows, ``ainv`` / ``ic`` and CHOLMOD come from pymathprim's CPU code, which is
not part of this path); hydra / lightning are replaced by plain arguments: ``run_inference(model, samples, ...)`` takes
any iterable of ``Data`` samples (``neural_cg.dataset.FolderDataset``, ``systems.*_sample``), and ``main`` is an
argparse front end over a dataset folder or a synthetic family.
"""
from __future__ import annotations

import argparse
import math
from dataclasses import dataclass
from pathlib import Path
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np
import pandas as pd
import torch

from .neural_cg.utils.validate import get_cg_iter_time, get_pcg_iter_time, get_pcg_scaled_iter_time, to_csr_cpu, to_numpy


@dataclass
class InferenceTimestat:
    all_solve_time: List[float]
    all_prec_time: List[float]
    all_iteration: List[float]
    all_matrix_size: List[int]


class Timestat:
    """Per-key lists of (solve s, preconditioner s, iterations, matrix size) and the two tables of infer.py:38-151."""

    SUMMARY_COLS = ["Total Time (ms)", "Solve Time (ms)", "Precond Time (ms)", "#Iteration"]
    ALL_COLS = ["Solve Time (ms)", "Precond Time (ms)", "#Iteration", "Matrix Size"]

    def __init__(self):
        self.stat_dict: Dict[str, InferenceTimestat] = {}

    def put(self, key: str, solve_time: float, prec_time: float, iteration: float, matrix_size: int):
        st = self.stat_dict.setdefault(key, InferenceTimestat([], [], [], []))
        st.all_solve_time.append(solve_time)
        st.all_prec_time.append(prec_time)
        st.all_iteration.append(iteration)
        st.all_matrix_size.append(matrix_size)

    def _means(self):
        for key, st in self.stat_dict.items():
            yield key, np.mean(st.all_solve_time) * 1000, np.mean(st.all_prec_time) * 1000, np.mean(st.all_iteration)

    def timestat_to_dataframe(self) -> pd.DataFrame:
        rows = [{"Key": k, "Total Time (ms)": s + p, "Solve Time (ms)": s, "Precond Time (ms)": p, "#Iteration": it}
                for k, s, p, it in self._means()]
        df = pd.DataFrame(rows, columns=["Key"] + self.SUMMARY_COLS)
        df[self.SUMMARY_COLS] = df[self.SUMMARY_COLS].astype(float).round(4)
        return df

    def all_time_stat(self) -> pd.DataFrame:
        rows = []
        for key, st in self.stat_dict.items():
            for s, p, i, m in zip(st.all_solve_time, st.all_prec_time, st.all_iteration, st.all_matrix_size):
                rows.append({"Key": key, "Solve Time (ms)": s * 1000, "Precond Time (ms)": p * 1000, "#Iteration": i,
                             "Matrix Size": m})
        df = pd.DataFrame(rows, columns=["Key"] + self.ALL_COLS)
        df[self.ALL_COLS] = df[self.ALL_COLS].astype(float).round(4)
        return df

    def print(self, out=print):
        for k, s, p, it in self._means():
            out(f"{k}: Average Solve Time: {s:.2f} ms Precond Time: {p:.2f} ms Iteration Time: {it:.4f} it/sample")

    def rich_print(self):
        try:
            from rich.console import Console
            from rich.table import Table
        except ImportError:
            return self.print()
        table = Table(title="Inference Time Statistics")
        table.add_column("Key", justify="left")
        for c in ("Total Time (ms)", "Solve Time (ms)", "Precond Time (ms)", "#Iteration"):
            table.add_column(c, justify="right")
        for k, s, p, it in self._means():
            table.add_row(k, f"{s + p:.2f}", f"{s:.2f}", f"{p:.2f}", f"{it:.4f}")
        Console().print(table)


def get_workspace(name: str):
    """(workspace class, neural PCG timing function) -- infer.py:153-163."""
    if name == "simple":
        from .neural_cg.workspace import SimpleTrainingWorkspace

        return SimpleTrainingWorkspace, get_pcg_iter_time
    if name == "scaled":
        from .neural_cg.scaled_workspace import ScaledTrainingWorkspace

        return ScaledTrainingWorkspace, get_pcg_scaled_iter_time
    raise ValueError(f"Unknown workspace name: {name}")


def make_rhs(kind: str, mask, A_full=None, rng: Optional[np.random.Generator] = None) -> np.ndarray:
    """The vector handed to the solvers as ``gt`` (they solve A x = A @ gt), infer.py:296-307."""
    m = to_numpy(mask).flatten().astype(np.float64)
    if kind in ("mask", "ones"):
        return m
    if kind == "random":
        r = (rng or np.random.default_rng()).standard_normal(m.size)
        return r * m
    if kind == "neighbour":
        if A_full is None:
            raise ValueError("rhs='neighbour' needs the unmasked pattern matrix")
        return (A_full @ (1 - m) + 0.1 * m) * m
    raise ValueError(f"Unknown rhs type: {kind}")


def _mat_size(sample, block_size: int) -> int:
    ptr = getattr(sample, "ptr", None)
    nodes = int(ptr[-1]) if ptr is not None else int(sample.x.shape[0])
    return nodes * block_size


def run_inference(model, samples: Iterable, pcg=get_pcg_iter_time, rhs: str = "mask", rtol: float = 1e-6, repeat: int = 1,
                  warmup: int = 20, methods: Sequence[str] = ("none", "diagonal"), device: str = "cuda",
                  seed: Optional[int] = None, progress=None) -> Timestat:
    """The sample loop of infer.py:262-331 on the CUDA path.  ``model`` is a workspace (``inference_step``, ``epsilon``,
    ``block_size``), ``samples`` yields ``Data`` objects (batch size 1, like the reference's inference loaders)."""
    if device != "cuda":
        raise ValueError(f"device={device!r}: the B200 path solves on 'cuda' only")
    stats = Timestat()
    rng = np.random.default_rng(seed)
    samples = iter(samples)
    first = next(samples, None)
    if first is None:
        return stats
    first = first.to(device)
    with torch.no_grad():
        for _ in range(max(int(warmup), 0)):       # warm up the GNN kernels / allocator (infer.py:270-275)
            model.inference_step(first)
        sample = first
        while sample is not None:
            sample = sample.to(device)
            n = _mat_size(sample, model.block_size)
            A = to_csr_cpu(sample.edge_index, sample.matrix_values, n, sample.mask)
            A_full = None
            if rhs == "neighbour":
                A_full = to_csr_cpu(sample.edge_index, sample.matrix_values, n, None)
                A_full.data.fill(1.0)
            prec = 0.0
            for _ in range(repeat):                  # preconditioner time = GNN forward + D2H (workspace.py:195-205)
                _, this = model.inference_step(sample)
                prec += this
            prec /= repeat
            L, _ = model.inference_step(sample)
            r = make_rhs(rhs, sample.mask, A_full, rng)
            msize = A.shape[0]
            try:
                for m in methods:
                    it, p, sol = get_cg_iter_time(A, r, rtol=rtol, repeat=repeat, method=m, device=device)
                    stats.put(f"PCG-{m}-{device}", sol, p, it, msize)
                it, _, sol = pcg(A, r, L, model.epsilon, rtol=rtol, device=device, repeat=repeat)
                stats.put("Neural+CUDA", sol, prec, it, msize)
            except RuntimeError as e:                # infer.py:362-363: a failed sample is logged, the run goes on
                print(f"RuntimeError: {e}")
            if progress is not None:
                progress(msize)
            sample = next(samples, None)
    return stats


def save_stats(stats: Timestat, out_dir, exp_name: str, rtol: float, infer_prefix: str = ""):
    """Writes the two CSV files of infer.py:372-384 and returns their paths."""
    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    log_rtol = -int(math.log10(rtol))
    fname = out_dir / f"infer_{infer_prefix}{exp_name}_{log_rtol}.csv"
    stats.timestat_to_dataframe().to_csv(fname, index=False)
    all_fname = out_dir / f"all_infer_{infer_prefix}{exp_name}_{log_rtol}.csv"
    stats.all_time_stat().to_csv(all_fname, index=False)
    return fname, all_fname


def main_hydra(argv: Sequence[str]):
    """``python -m learningsparsepreconditioner4gpu_b200.infer exp_name=heat pretrained=ckpt.pt +rtol=1e-6 ...``: the
    reference's Hydra entry point (infer.py:166-385) on the composed ``config/*.yaml`` -- same keys, same ad-hoc
    ``cfg.get`` knobs (workspace, repeat, rhs, rtol, warmup, dataloader, out_dir, infer_prefix; infer.py:218-259,274,376).
    ``pretrained`` is a plain ``state_dict`` file (Lightning ``.ckpt`` pickles need omegaconf + lightning); empty ->
    the seed-42 random init."""
    from .neural_cg.utils import config as hc
    from .neural_cg.utils.datamodule import FolderDataModule

    name, cdir, overrides = hc.parse_cli(list(argv))
    cfg = hc.compose(name, overrides, cdir)
    if "all_prefix" in cfg.data:
        raise NotImplementedError("MultiFolderDataModule (data.all_prefix) is outside the solve path")
    data = FolderDataModule(data_config=cfg.data, split_config=cfg.split, batch_size=cfg.batch_size)
    edge_f, node_f, block_size = data.dataset.num_edge_features, data.dataset.num_node_features, data.dataset.block_size
    workspace_class, pcg = get_workspace(cfg.get("workspace", "simple"))
    kw = hc.workspace_kwargs(cfg)
    torch.manual_seed(int(cfg.get("seed", 42)))
    model = workspace_class(edge_features=edge_f, node_features=node_f, block_size=block_size, **kw).eval()
    if cfg.get("pretrained", ""):
        sd = torch.load(cfg.pretrained, map_location="cpu")
        model.load_state_dict(sd.get("state_dict", sd))
    loader_name = cfg.get("dataloader", "test")
    if loader_name == "train":
        dl = data.train_dataloader()
    elif loader_name in ("val", "test"):
        dl = data.val_dataloader()
    elif loader_name == "infer":
        dl = data.inference_dataloader()
    else:
        raise ValueError(f"Unknown dataloader name: {loader_name}")
    rtol = float(cfg.get("rtol", 1e-6))
    stats = run_inference(model, dl, pcg=pcg, rhs=cfg.get("rhs", "mask"), rtol=rtol, repeat=int(cfg.get("repeat", 1)),
                          warmup=int(cfg.get("warmup", 20)))
    stats.rich_print()
    f1, f2 = save_stats(stats, cfg.get("out_dir", "output"), cfg.exp_name, rtol, cfg.get("infer_prefix", ""))
    print(f"Inference statistics saved to {f1}\nAll inference statistics saved to {f2}")
    return stats


def main(argv=None):
    import sys

    argv = list(sys.argv[1:] if argv is None else argv)
    # Hydra-style command line (key=value / +key=value overrides, as the reference is launched) -> the config path;
    # the --flag form below stays as an alias for synthetic systems and ad-hoc folders
    if any((not a.startswith("--")) and ("=" in a or a.startswith(("+", "~"))) for a in argv):
        return main_hydra(argv)
    ap = argparse.ArgumentParser(description="learned-SPAI PCG inference benchmark (reference infer.py) on B200")
    ap.add_argument("--exp-name", default="poisson3d")
    ap.add_argument("--data-prefix", default=None, help="dataset folder (reference layout); default: synthetic systems")
    ap.add_argument("--fixed-topology", action="store_true")
    ap.add_argument("--block-size", type=int, default=1)
    ap.add_argument("--systems", type=int, default=4, help="synthetic: number of systems")
    ap.add_argument("--grid", type=int, default=24, help="synthetic: grid nodes per axis")
    ap.add_argument("--workspace", default="simple", choices=["simple", "scaled"])
    ap.add_argument("--pretrained", default=None, help="state_dict checkpoint (torch.save) of the workspace")
    ap.add_argument("--rhs", default="mask", choices=["mask", "ones", "random", "neighbour"])
    ap.add_argument("--rtol", type=float, default=1e-6)
    ap.add_argument("--repeat", type=int, default=1)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--out-dir", default="output")
    ap.add_argument("--infer-prefix", default="")
    args = ap.parse_args(argv)
    from . import systems

    if args.data_prefix:
        from .neural_cg.dataset import FolderDataset

        ds = FolderDataset(is_fixed_topology=args.fixed_topology, load_into_memory=False, block_size=args.block_size,
                           has_shared_features=(Path(args.data_prefix) / "shared_features.npy").exists(),
                           use_node_features=any((Path(args.data_prefix) / "features").glob("*.npy")),
                           use_matrix_as_edge_feature=True, use_mask_as_node_feature=True,
                           use_node_features_as_edge_feature=False, use_edge_features_as_node_feature="disable",
                           use_random_rhs=True, normalize_matrix="mean", prefix=args.data_prefix)
        samples = [ds[i] for i in range(len(ds))]
    else:
        samples = [systems.poisson_sample(args.grid, seed=100 + k) for k in range(args.systems)]
    _, pcg = get_workspace(args.workspace)
    model = systems.default_workspace(samples[0], block_size=args.block_size, scaled=args.workspace == "scaled")
    if args.pretrained:
        sd = torch.load(args.pretrained, map_location="cpu")
        model.load_state_dict(sd.get("state_dict", sd))
    stats = run_inference(model, samples, pcg=pcg, rhs=args.rhs, rtol=args.rtol, repeat=args.repeat, warmup=args.warmup)
    stats.rich_print()
    f1, f2 = save_stats(stats, args.out_dir, args.exp_name, args.rtol, args.infer_prefix)
    print(f"Inference statistics saved to {f1}\nAll inference statistics saved to {f2}")


if __name__ == "__main__":
    main()
