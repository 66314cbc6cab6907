"""Host-side mirror of ``neural_cg/workspace.py``: ``SimpleTrainingWorkspace`` with the reference's
constructor signature (:27-50), attributes (``gnn``, ``epsilon``, ``block_size``, ``spai_preconditioner``)
and inference methods (``forward`` :92-94, ``inference_step`` :195-205, ``precondition`` :207-211,
``log_converge_batch`` :114-143, ``log_converge_v2`` :145-166).

Differences, all forced by the environment or the scope:
* a plain ``nn.Module`` (Lightning is absent): ``log`` / ``log_dict`` collect into ``self.logged``,
  ``save_hyperparameters`` fills ``self.hparams``; checkpoints are plain ``state_dict``s with the reference's keys;
* the GNN forward, CSR assembly and PCG run in libb200spai.so; convergence checks always solve on "cuda"
  (a "cpu" entry in ``check_devices`` is mapped to "cuda" with a warning) with the methods the library has;
* training: ``_shared_step`` / ``training_step`` / ``validation_step`` / ``test_step`` /
  ``configure_optimizers`` (:96-112,178-193,213-226) run the GNN forward with a tape, ``AATPE`` and the loss
  (``neural_cg/loss.py``) on the device; ``loss.backward()`` runs the CUDA backward kernels (``b200spai_gnn_backward``,
  ``b200spai_graph_outer_f32``) behind torch autograd Functions.  ``neural_cg/trainer.py`` is the small stand-in for the
  ``L.Trainer(...).fit`` loop of train.py:98-110;
* additions: ``inference_step_device`` (G stays in HBM) and ``solve`` (GNN -> G -> PCG, end to end on the device).
"""
from __future__ import annotations

import warnings
from time import time
from typing import Any, Dict, List, Optional, Union

import numpy as np
import torch
from scipy.sparse import csr_matrix
from torch import nn

from .._lib import nvtx_range
from ..device import DeviceCSR, DevicePCG
from .loss import create_loss_item
from .nn.basic_layers import AATPE
from .nn.gnns import NodeEdgeProcessing
from .utils.optim import create_optimizer, create_scheduler
from .utils.validate import get_cg_iter_time, get_pcg_iter_time, to_csr_cpu, to_csr_device, to_numpy


def _cfg_get(cfg, key, default=None):
    if cfg is None:
        return default
    if isinstance(cfg, dict):
        return cfg.get(key, default)
    return getattr(cfg, key, default)


class FixedTopology:
    """Caller-owned cache of everything that depends only on (edge_index, mask) -- the reference's default dataset mode
    (config/data.yaml ``is_fixed_topology: True``: one edge list for all samples): the GNN's target-sorted permutation and
    segment offsets, the CSR patterns of A and G with their row-block partitions and compressed column streams, the
    permutation of G^T, the solver's workspaces and captured graphs.  ``workspace.solve(sample, topology=cache)`` fills it
    on the first sample and afterwards only moves VALUES (one gather pass per matrix).  Every use checks that the sample's
    edge_index and mask still equal the cached ones (device compare, 0.4 ms at 150 M edges) and that no exact zero
    appeared or disappeared among the values (the reference drops exact zeros from the pattern); either rebuilds."""

    def __init__(self):
        self.edge_index = self.mask = None
        self.graph = self.A = self.G = self.solver = None
        self.rebuilds = 0

    def matches(self, sample) -> bool:
        return (self.edge_index is not None and self.edge_index.shape == sample.edge_index.shape
                and self.mask.shape == sample.mask.shape and self.edge_index.device == sample.edge_index.device
                and bool(torch.equal(self.edge_index, sample.edge_index)) and bool(torch.equal(self.mask, sample.mask)))

    def clear(self):
        for h in (self.solver, self.A, self.G, self.graph):
            if h is not None:
                h.close()
        self.__init__()


class _WorkspaceBase(nn.Module):
    """What the two workspaces share: construction, logging stand-ins, the per-sample convergence probe."""

    def __init__(self, batch_size: int, batch_less: bool, block_size: int, test_max_iter: int, optimizer, scheduler, loss,
                 inspect_norms: bool, check_converge: bool, check_methods: List[str], check_devices: List[str],
                 node_features: int, edge_features: int, gnn, epsilon: float, *args, **kwargs):
        super().__init__()
        self.hparams: Dict[str, Any] = dict(
            batch_size=batch_size, batch_less=batch_less, block_size=block_size, test_max_iter=test_max_iter,
            optimizer=optimizer, scheduler=scheduler, loss=loss, inspect_norms=inspect_norms, check_converge=check_converge,
            check_methods=check_methods, check_devices=check_devices, node_features=node_features,
            edge_features=edge_features, gnn=gnn, epsilon=epsilon, **kwargs)
        self.block_size = block_size
        self.optimizer_cfg = optimizer
        self.scheduler_cfg = scheduler
        self.epsilon = epsilon
        self.test_max_iter = test_max_iter
        self.loss_name = _cfg_get(loss, "name")
        self.loss_fn = create_loss_item(self.loss_name, batch_less=batch_less, block_size=block_size,
                                        **dict(_cfg_get(loss, "params") or {})) if self.loss_name else None
        self.gnn = NodeEdgeProcessing(node_in_features=node_features, node_out_features=None,
                                      edge_in_features=edge_features, edge_out_features=block_size * block_size,
                                      **dict(gnn))
        self.enable_inspect_norms = inspect_norms
        self.spai_preconditioner = AATPE(self.epsilon)
        self.check_converge = check_converge
        self.check_methods: List[str] = list(check_methods or [])
        self.check_devices: List[str] = list(check_devices or [])
        self.has_pymathprim = False  # the solver is libb200spai; kept for callers that inspect it
        self.logged: Dict[str, Any] = {}

    # ---- Lightning stand-ins -----------------------------------------------------------------------------
    def log(self, name: str, value, **kwargs) -> None:
        self.logged[name] = float(value)

    def log_dict(self, values: Dict[str, Any], **kwargs) -> None:
        for k, v in values.items():
            self.log(k, v)

    # ---- training (workspace.py:96-112,178-193,213-226) ---------------------------------------------------------
    def _precondition_residual(self, batch, boo_entry):
        return self.spai_preconditioner(batch.residual, batch.edge_index, boo_entry, mask=batch.mask)

    def _shared_step(self, batch, prefix: str, batch_idx: int):
        assert hasattr(batch, "ptr"), "Relying on batch.ptr to compute batch size"
        if self.loss_fn is None:
            raise ValueError("the workspace was built without a loss config")
        bsize = int(batch.ptr.shape[0]) - 1
        with nvtx_range("b200spai/train_forward"):
            boo_entry = self(batch.x, batch.edge_index, batch.edge_attr, autograd=True)
        with nvtx_range("b200spai/train_aatpe+loss"):
            d = self._precondition_residual(batch, boo_entry)
            total_loss = self.loss_fn(batch, d, boo_entry)
        self.log_dict({f"{prefix}/Loss": total_loss.detach()}, batch_size=bsize, prog_bar=(prefix == "Train"))
        # (the reference probes convergence here for Val / Test at batch 0 AND again in validation_step / test_step,
        # workspace.py:107-108,184-185: the same numbers twice; _eval_step below runs the probe once)
        self._log_norms(boo_entry.detach(), bsize, prefix)
        return boo_entry, total_loss

    def training_step(self, batch, batch_idx: int):
        _, loss = self._shared_step(batch, "Train", batch_idx)
        return loss

    def _eval_step(self, batch, prefix: str, batch_idx: int, repeat: int, every_batch: bool):
        """validation_step / test_step (workspace.py:182-193): loss + convergence probe.  Batches made for inference only
        (no ``residual``: ``make_data(is_inference=True)``) have no loss; they get the probe and return G."""
        if "residual" in batch and self.loss_fn is not None:
            boo_entry, out = self._shared_step(batch, prefix, batch_idx)
            boo_entry = boo_entry.detach()
        else:
            with torch.no_grad():
                boo_entry = out = self(batch.x, batch.edge_index, batch.edge_attr)
            self._log_norms(boo_entry, int(batch.ptr.shape[0]) - 1, prefix)
        if every_batch or batch_idx == 0:
            self.log_converge_batch(boo_entry, batch, prefix, repeat)
        return out

    def validation_step(self, batch, batch_idx: int):
        return self._eval_step(batch, "Val", batch_idx, repeat=1, every_batch=False)

    def configure_optimizers(self):
        opt = create_optimizer(self.parameters(), _cfg_get(self.optimizer_cfg, "name"), _cfg_get(self.optimizer_cfg, "params"))
        sched = create_scheduler(opt, _cfg_get(self.scheduler_cfg, "name"), _cfg_get(self.scheduler_cfg, "params"))
        return {"optimizer": opt, "lr_scheduler": sched} if sched is not None else opt

    # ---- the model -----------------------------------------------------------------------------------------
    def forward(self, node_attr, edge_index, edge_attr, autograd: bool = False):
        """[nE, b, b].  ``autograd=True`` (the training / validation steps): the result carries the grad_fn of the CUDA
        backward; every other caller gets the plain inference forward."""
        _, boo_entry = self.gnn(node_attr, edge_index, edge_attr, autograd=autograd)
        return boo_entry.reshape(-1, self.block_size, self.block_size)  # [nE, b, b]

    def _devices(self) -> List[str]:
        out = []
        for d in self.check_devices:
            if d != "cuda":
                warnings.warn(f"check_devices entry {d!r} is solved on 'cuda': this build has no CPU solver", stacklevel=3)
            if "cuda" not in out:
                out.append("cuda")
        return out

    def _pcg(self, A, r, L, device, repeat):
        return get_pcg_iter_time(A, r, L, self.epsilon, device=device, repeat=repeat)

    def log_converge_v2(self, L, A, r, repeat, prefix: str):
        stats: Dict[str, float] = {}
        for device in self._devices():
            for method in self.check_methods:
                if method not in ("none", "diagonal"):
                    continue  # ic / ainv / fsai: reference comparison baselines (pymathprim), not in this path
                avg_iter, _, avg_solve = get_cg_iter_time(A, r, repeat=repeat, method=method, device=device)
                stats[f"{prefix}/{device}_{method}_iter"] = stats.get(f"{prefix}/{device}_{method}_iter", 0) + avg_iter
                stats[f"{prefix}/{device}_{method}_solve"] = stats.get(f"{prefix}/{device}_{method}_solve", 0) + avg_solve
            avg_iter, _, avg_solve = self._pcg(A, r, L, device, repeat)
            stats[f"{prefix}/{device}_neural_iter"] = avg_iter
            stats[f"{prefix}/{device}_neural_solve"] = avg_solve
        self.log_dict(stats, batch_size=1)
        return stats

    def _export_factor(self, L: csr_matrix, batch) -> csr_matrix:
        return L

    def log_converge_batch(self, boo_entry: torch.Tensor, batch, prefix: str, repeat=1):
        """Split the block-diagonal batch by ``ptr`` and solve every sample on its own (workspace.py:137-143)."""
        if not self.check_converge:
            return None
        ptr = batch.ptr.detach().cpu().numpy().tolist()
        n = ptr[-1] * self.block_size
        L = self._export_factor(to_csr_cpu(batch.edge_index, boo_entry, n, batch.mask, dtype=np.float64), batch)
        A = to_csr_cpu(batch.edge_index, batch.matrix_values, n, batch.mask, dtype=np.float64)
        r = to_numpy(batch.mask).flatten().astype(np.float64, copy=True)
        all_stats = []
        for b, e in zip(ptr[:-1], ptr[1:]):
            b, e = b * self.block_size, e * self.block_size
            all_stats.append(self.log_converge_v2(csr_matrix(L[b:e, b:e]), csr_matrix(A[b:e, b:e]), r[b:e], repeat, prefix))
        return all_stats

    def _log_norms(self, boo_entry: torch.Tensor, bsize: int, prefix: str):
        if self.enable_inspect_norms:
            self.log(f"{prefix}/boo_entry_norm", torch.linalg.norm(boo_entry), batch_size=1)

    # ---- device-resident additions ---------------------------------------------------------------------------
    @torch.no_grad()
    def inference_step_device(self, sample, pattern: str = "full") -> DeviceCSR:
        """G as a DeviceCSR without leaving HBM (GNN forward + device assembly).

        ``pattern="full"`` is the reference's contract on this path: one block per stored entry of A (workspace.py:92-94).
        ``pattern="tril"`` keeps the blocks with col <= row -- ``ToLowerTriangular`` (basic_layers.py:278-294) applied
        to the GNN output before assembly -- so G is (block) lower triangular and M = G G^T + eps I stays SPD."""
        boo_entry = self(sample.x, sample.edge_index, sample.edge_attr)
        edge_index = sample.edge_index
        if pattern == "tril":
            from .nn.basic_layers import ToLowerTriangular

            boo_entry, edge_index = ToLowerTriangular()(boo_entry, edge_index)
            boo_entry, edge_index = boo_entry.contiguous(), edge_index.contiguous()
        elif pattern != "full":
            raise ValueError(f"pattern must be 'full' or 'tril', not {pattern!r}")
        n = sample.x.shape[0] * self.block_size
        return to_csr_device(edge_index, boo_entry, n, sample.mask)

    @torch.no_grad()
    def solve(self, sample, rhs: Optional[torch.Tensor] = None, rtol: float = 1e-6, max_iter: int = 0,
              preconditioner: str = "ext_spai", history: bool = False, topology: Optional[FixedTopology] = None,
              pattern: str = "full"):
        """GNN -> G -> PCG on the device.  ``rhs`` defaults to the reference's rhs="mask" convention
        (b = A @ mask, infer.py:297-299).  Returns (x, PcgResult, history, A, G).  ``topology``: a ``FixedTopology`` cache
        shared by the samples of a fixed-topology dataset (A, G and the solver then belong to the cache).  ``pattern``:
        "full" (the reference's contract) or "tril" (see ``inference_step_device``)."""
        if topology is not None:
            if pattern != "full":
                raise ValueError("the fixed-topology cache keeps the reference's full pattern of G")
            return self._solve_fixed(sample, topology, rhs, rtol, max_iter, preconditioner, history)
        n = sample.x.shape[0] * self.block_size
        with nvtx_range("b200spai/assemble_A"):
            A = to_csr_device(sample.edge_index, sample.matrix_values, n, sample.mask)
        with nvtx_range("b200spai/gnn+assemble_G"):
            G = self.inference_step_device(sample, pattern)
        if rhs is None:
            rhs = A.matvec(sample.mask.reshape(-1).to(torch.float64))
        with nvtx_range("b200spai/pcg_setup"):
            solver = DevicePCG(A)
            solver.set_spai(G, self.epsilon)
        with nvtx_range("b200spai/pcg_solve"):
            x, res, hist = solver.solve(rhs.to(torch.float64).reshape(-1), preconditioner, rtol=rtol, max_iter=max_iter,
                                        history=history)
        return x, res, hist, A, G


    def _solve_fixed(self, sample, topo: FixedTopology, rhs, rtol, max_iter, preconditioner, history):
        from .nn.gnns import DeviceGraph

        N = sample.x.shape[0]
        n = N * self.block_size
        fresh = not topo.matches(sample)
        if fresh:
            topo.clear()
            topo.edge_index, topo.mask = sample.edge_index.clone(), sample.mask.clone()
            topo.graph = DeviceGraph(topo.edge_index, N)
        _, g_vals = self.gnn(sample.x, topo.edge_index, sample.edge_attr, graph=topo.graph)
        g_vals = g_vals.reshape(-1, self.block_size, self.block_size)
        if not fresh:
            ok = topo.A.refresh_from_edges(sample.matrix_values) and topo.G.refresh_from_edges(g_vals)
            if ok:
                topo.solver.refresh()
            else:                       # an exact zero appeared or disappeared: the reference's pattern changed
                fresh = True
                topo.rebuilds += 1
                for h in (topo.solver, topo.A, topo.G):
                    h.close()
        if fresh:
            topo.A = to_csr_device(topo.edge_index, sample.matrix_values, n, topo.mask, plan=True)
            topo.G = to_csr_device(topo.edge_index, g_vals, n, topo.mask, plan=True)
            topo.solver = DevicePCG(topo.A)
            topo.solver.set_spai(topo.G, self.epsilon)
        if rhs is None:
            rhs = topo.A.matvec(sample.mask.reshape(-1).to(torch.float64))
        x, res, hist = topo.solver.solve(rhs.to(torch.float64).reshape(-1), preconditioner, rtol=rtol, max_iter=max_iter,
                                         history=history)
        return x, res, hist, topo.A, topo.G


class SimpleTrainingWorkspace(_WorkspaceBase):
    @torch.no_grad()
    def inference_step(self, sample, time_beg: Union[float, None] = None):
        time_beg = time()
        boo_entry = self(sample.x, sample.edge_index, sample.edge_attr).cpu().detach().numpy()
        dt = time() - time_beg  # the preconditioner-construction time, including the D2H copy (workspace.py:200-202)
        n = sample.x.shape[0] * self.block_size
        csr = to_csr_cpu(sample.edge_index, boo_entry, n, sample.mask, dtype=np.float64)
        return csr, dt

    def precondition(self, sample, Adense) -> np.ndarray:
        L, _ = self.inference_step(sample)
        M = L @ L.T
        return M.todense() + self.epsilon * np.eye(M.shape[0], dtype=np.float64)

    def test_step(self, batch, batch_idx: int):
        out = self._eval_step(batch, "Test", batch_idx, repeat=3, every_batch=True)
        _, dt = self.inference_step(batch)
        self.log("Test/time_neural_prec", dt, batch_size=1)
        return out
