"""Host-side mirror of ``neural_cg/scaled_workspace.py``: the Jacobi-scaled variant
M^-1 = G D^-1 G^T + eps D^-1 (:101).  At export the factor is right-multiplied by diag(rsqrt_diag)
(:138, :210-211) and the solve uses the "ext_spai_scaled" preconditioner (validate.py:124-160).

Note carried over from the reference: its scipy definition of "ext_spai_scaled" divides by
diag(A) again, so exported-and-scaled G is scaled twice relative to the training operator.  The kernels
implement the reference's executable definition (validate.py:276-287) as is.
"""
from __future__ import annotations

from time import time
from typing import Union

import numpy as np
import torch
from scipy.sparse import csr_matrix, diags

from .utils.validate import get_pcg_scaled_iter_time, to_csr_cpu, to_numpy
from .workspace import _WorkspaceBase


class ScaledTrainingWorkspace(_WorkspaceBase):
    def _pcg(self, A, r, L, device, repeat):
        return get_pcg_scaled_iter_time(A, r, L, self.epsilon, device=device, repeat=repeat)

    def _precondition_residual(self, batch, boo_entry):
        assert hasattr(batch, "inv_diag"), "Relying on batch.inv_diag to maintain stability"   # scaled_workspace.py:98
        return self.spai_preconditioner(batch.residual, batch.edge_index, boo_entry, mask=batch.mask, diag=batch.inv_diag)

    def _export_factor(self, L: csr_matrix, batch) -> csr_matrix:
        rsqrt_diag = to_numpy(batch.rsqrt_diag).flatten().astype(np.float64, copy=True)
        return csr_matrix(L @ diags(rsqrt_diag))

    @torch.no_grad()
    def inference_step(self, sample, time_beg: Union[float, None] = None):
        time_beg = time_beg or time()
        boo_entry = self(sample.x, sample.edge_index, sample.edge_attr)
        torch.cuda.current_stream().synchronize()  # the reference reads the clock before any sync (:205-207)
        dt = time() - time_beg
        n = sample.x.shape[0] * self.block_size
        csr = to_csr_cpu(sample.edge_index, boo_entry, n, sample.mask, dtype=np.float64)
        return self._export_factor(csr, sample), dt

    def precondition(self, sample, Adense) -> np.ndarray:
        L, _ = self.inference_step(sample)
        dinv = np.diag(1.0 / np.diag(Adense))
        M = L @ dinv @ L.T + self.epsilon * dinv
        return np.asarray(M) + self.epsilon * np.eye(M.shape[0], dtype=np.float64)
