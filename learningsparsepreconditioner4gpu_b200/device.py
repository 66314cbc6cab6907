"""Thin Python handles over the C ABI: device CSR matrices and the PCG solver.

PyTorch is used only for device memory (tensors passed as raw pointers), the current stream and
synchronisation.  Every arithmetic step happens in libb200spai.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib

PRECOND = {"none": 0, "diagonal": 1, "ext_spai": 2, "ext_spai_scaled": 3}


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t) -> C.c_void_p:
    if t is None:
        return C.c_void_p(0)
    if isinstance(t, torch.Tensor):
        assert t.is_contiguous()
        return C.c_void_p(t.data_ptr())
    if isinstance(t, np.ndarray):
        assert t.flags["C_CONTIGUOUS"]
        return C.c_void_p(t.ctypes.data)
    raise TypeError(type(t))


class DeviceCSR:
    """fp64 CSR matrix resident in HBM (b200spai_csr_t)."""

    def __init__(self, handle: C.c_void_p):
        self._h = handle
        self._lib = _lib.load()

    # ---- constructors --------------------------------------------------------------------------
    @classmethod
    def from_scipy(cls, m) -> "DeviceCSR":
        from scipy.sparse import csr_matrix

        _lib.require_cuda()
        m = csr_matrix(m)
        if not m.has_sorted_indices:
            m = m.sorted_indices()
        indptr = np.ascontiguousarray(m.indptr, dtype=np.int32)
        indices = np.ascontiguousarray(m.indices, dtype=np.int32)
        data = np.ascontiguousarray(m.data, dtype=np.float64)
        h = C.c_void_p()
        lib = _lib.load()
        _lib.check(lib.b200spai_csr_create_host(m.shape[0], m.shape[1], m.nnz, _ptr(indptr), _ptr(indices), _ptr(data),
                                                _stream(), C.byref(h)), "csr_create_host")
        return cls(h)

    @classmethod
    def from_torch(cls, indptr: torch.Tensor, indices: torch.Tensor, data: torch.Tensor, shape) -> "DeviceCSR":
        assert indptr.is_cuda and indptr.dtype == torch.int32 and indices.dtype == torch.int32 and data.dtype == torch.float64
        h = C.c_void_p()
        lib = _lib.load()
        _lib.check(lib.b200spai_csr_create_dev(shape[0], shape[1], indices.numel(), _ptr(indptr.contiguous()),
                                               _ptr(indices.contiguous()), _ptr(data.contiguous()), _stream(), C.byref(h)),
                   "csr_create_dev")
        torch.cuda.current_stream().synchronize()
        return cls(h)

    @classmethod
    def from_edges(cls, edge_index: torch.Tensor, values: torch.Tensor, num_nodes: int,
                   mask: Optional[torch.Tensor] = None, plan: bool = False) -> "DeviceCSR":
        """Device to_csr_cpu (validate.py:22-51): values fp32 [E] / [E,1,1] / [E,b,b], mask fp32 [N,b] or None.
        ``plan``: also record where every stored value comes from, so that ``refresh_from_edges`` can re-fill the matrix
        from the next sample's values of the SAME topology and mask in one gather pass (fixed-topology datasets)."""
        _lib.require_cuda()
        assert edge_index.is_cuda and edge_index.dtype == torch.int64 and edge_index.shape[0] == 2
        b = 1 if values.dim() == 1 else values.shape[-1]
        vals = values.detach().to(torch.float32).contiguous()
        m = None if mask is None else mask.detach().to(torch.float32).contiguous()
        ei = edge_index.contiguous()
        h = C.c_void_p()
        lib = _lib.load()
        fn = lib.b200spai_edges_to_csr_plan if plan else lib.b200spai_edges_to_csr
        _lib.check(fn(num_nodes, ei.shape[1], b, _ptr(ei), _ptr(vals), _ptr(m), _stream(), C.byref(h)), "edges_to_csr")
        return cls(h)

    @property
    def plan_kind(self) -> int:
        """0 = no refresh plan, 1 = planned from edges, 2 = planned transpose."""
        k = C.c_int32()
        _lib.check(self._lib.b200spai_csr_has_plan(self._h, C.byref(k)), "csr_has_plan")
        return k.value

    def refresh_from_edges(self, values: torch.Tensor) -> bool:
        """New values on the planned pattern.  False: an exact zero appeared / disappeared, the reference's pattern would
        change -- the content is undefined, rebuild with ``from_edges``."""
        vals = values.detach().to(torch.float32).contiguous()
        ok = C.c_int32()
        _lib.check(self._lib.b200spai_csr_refresh_from_edges(self._h, _ptr(vals), vals.numel(), _stream(), C.byref(ok)),
                   "csr_refresh_from_edges")
        return bool(ok.value)

    # ---- queries ---------------------------------------------------------------------------------
    @property
    def shape(self) -> Tuple[int, int]:
        r, c, z = C.c_int32(), C.c_int32(), C.c_int64()
        _lib.check(self._lib.b200spai_csr_shape(self._h, C.byref(r), C.byref(c), C.byref(z)), "csr_shape")
        return (r.value, c.value)

    @property
    def nnz(self) -> int:
        r, c, z = C.c_int32(), C.c_int32(), C.c_int64()
        _lib.check(self._lib.b200spai_csr_shape(self._h, C.byref(r), C.byref(c), C.byref(z)), "csr_shape")
        return z.value

    def stream_bytes(self):
        """(value bytes, index bytes) per stored entry that the streaming SpMV reads (4/8, 2/4)."""
        v, i = C.c_int32(), C.c_int32()
        _lib.check(self._lib.b200spai_csr_stream_info(self._h, C.byref(v), C.byref(i), _stream()), "csr_stream_info")
        return v.value, i.value

    def bsr_info(self):
        """(block size, stored blocks) of the native block stream the streaming SpMV uses: (3, nblocks) for matrices
        assembled from 3x3 blocks when that stream is the smaller one, else (1, 0)."""
        b, nb = C.c_int32(), C.c_int64()
        _lib.check(self._lib.b200spai_csr_bsr_info(self._h, _stream(), C.byref(b), C.byref(nb)), "csr_bsr_info")
        return b.value, nb.value

    def to_scipy(self):
        from scipy.sparse import csr_matrix

        n, m = self.shape
        nnz = self.nnz
        indptr = np.empty(n + 1, dtype=np.int32)
        indices = np.empty(nnz, dtype=np.int32)
        data = np.empty(nnz, dtype=np.float64)
        _lib.check(self._lib.b200spai_csr_download(self._h, _ptr(indptr), _ptr(indices), _ptr(data), _stream()), "csr_download")
        return csr_matrix((data, indices, indptr), shape=(n, m))

    def transpose(self) -> "DeviceCSR":
        h = C.c_void_p()
        _lib.check(self._lib.b200spai_csr_transpose(self._h, _stream(), C.byref(h)), "csr_transpose")
        return DeviceCSR(h)

    def diagonal(self) -> torch.Tensor:
        d = torch.empty(self.shape[0], dtype=torch.float64, device="cuda")
        _lib.check(self._lib.b200spai_csr_diagonal(self._h, _ptr(d), _stream()), "csr_diagonal")
        return d

    def matvec(self, x: torch.Tensor, out: Optional[torch.Tensor] = None, kernel: int = 0) -> torch.Tensor:
        assert x.is_cuda and x.dtype == torch.float64 and x.numel() == self.shape[1]
        y = out if out is not None else torch.empty(self.shape[0], dtype=torch.float64, device=x.device)
        _lib.check(self._lib.b200spai_spmv(self._h, _ptr(x.contiguous()), _ptr(y), kernel, _stream()), "spmv")
        return y

    def close(self):
        if self._h:
            self._lib.b200spai_csr_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DevicePCG:
    """PCG solver bound to a DeviceCSR A (b200spai_pcg_t)."""

    def __init__(self, A: DeviceCSR, mode: int = 0):
        self._lib = _lib.load()
        self.A = A
        self.G = None
        self._h = C.c_void_p()
        _lib.check(self._lib.b200spai_pcg_create(A._h, C.byref(self._h)), "pcg_create")
        if mode:
            self.set_mode(mode)

    def set_mode(self, mode: int):
        _lib.check(self._lib.b200spai_pcg_set_mode(self._h, mode), "pcg_set_mode")

    def refresh(self):
        """A and / or G were refreshed in place (``DeviceCSR.refresh_from_edges``): re-gather G^T, drop the cached diagonal."""
        _lib.check(self._lib.b200spai_pcg_refresh(self._h, _stream()), "pcg_refresh")

    def set_spai(self, G: DeviceCSR, epsilon: float):
        self.G = G  # keep alive
        _lib.check(self._lib.b200spai_pcg_set_spai(self._h, G._h, float(epsilon), _stream()), "pcg_set_spai")

    def solve(self, b: torch.Tensor, precond: str = "ext_spai", rtol: float = 1e-6, max_iter: int = 0,
              x: Optional[torch.Tensor] = None, history: bool = False):
        """Device-resident solve. Returns (x, PcgResult, history or None)."""
        assert b.is_cuda and b.dtype == torch.float64
        n = self.A.shape[0]
        x = x if x is not None else torch.empty(n, dtype=torch.float64, device=b.device)
        hist = None
        if history:
            hist = torch.zeros((max_iter if max_iter > 0 else n) + 1, dtype=torch.float64, device=b.device)
        res = _lib.PcgResult()
        _lib.check(self._lib.b200spai_pcg_solve_dev(self._h, PRECOND[precond], _ptr(b.contiguous()), _ptr(x), float(rtol),
                                                    int(max_iter), _ptr(hist), _stream(), C.byref(res)), "pcg_solve_dev")
        if hist is not None:
            hist = hist[: res.iterations + 1]
        return x, res, hist

    def solve_batch(self, row_ptr, b: torch.Tensor, precond: str = "ext_spai", rtol: float = 1e-6, max_iter: int = 0,
                    x: Optional[torch.Tensor] = None, history: bool = False):
        """Many independent systems in one launch per wave (b200spai_pcg_solve_batch): A (and G) are the block-diagonal
        matrices of a batch, block i = rows [row_ptr[i], row_ptr[i+1]).  Returns (x, [PcgResult per system], histories or
        None) -- histories[i] is system i's ||r_k||, k = 0..iterations."""
        assert b.is_cuda and b.dtype == torch.float64
        ptr = np.ascontiguousarray(np.asarray(row_ptr, dtype=np.int64))
        n_sys = int(ptr.size) - 1
        n = self.A.shape[0]
        assert n_sys >= 1 and ptr[0] == 0 and ptr[-1] == n
        ptr32 = ptr.astype(np.int32)
        x = x if x is not None else torch.empty(n, dtype=torch.float64, device=b.device)
        hist, stride = None, 0
        if history:
            stride = (int(max_iter) if max_iter > 0 else int(np.diff(ptr).max())) + 1
            hist = torch.zeros(n_sys * stride, dtype=torch.float64, device=b.device)
        res = (_lib.PcgResult * n_sys)()
        _lib.check(self._lib.b200spai_pcg_solve_batch(self._h, PRECOND[precond], n_sys, ptr32.ctypes.data_as(C.POINTER(C.c_int32)),
                                                      _ptr(b.contiguous()), _ptr(x), float(rtol), int(max_iter), _ptr(hist),
                                                      int(stride), _stream(), res), "pcg_solve_batch")
        results = list(res)
        hists = None
        if hist is not None:
            hists = [hist[i * stride: i * stride + results[i].iterations + 1] for i in range(n_sys)]
        return x, results, hists

    def solve_host(self, b: np.ndarray, x: np.ndarray, precond: str = "ext_spai", rtol: float = 1e-6, max_iter: int = 0,
                   history: Optional[np.ndarray] = None):
        """Host-buffer solve (copies inside the call); x is updated in place."""
        assert b.dtype == np.float64 and x.dtype == np.float64
        res = _lib.PcgResult()
        _lib.check(self._lib.b200spai_pcg_solve_host(self._h, PRECOND[precond], _ptr(b), _ptr(x), float(rtol), int(max_iter),
                                                     _ptr(history), _stream(), C.byref(res)), "pcg_solve_host")
        return res

    def apply_precond(self, r: torch.Tensor, precond: str = "ext_spai") -> torch.Tensor:
        z = torch.empty_like(r)
        _lib.check(self._lib.b200spai_pcg_apply_precond(self._h, PRECOND[precond], _ptr(r.contiguous()), _ptr(z), _stream()),
                   "pcg_apply_precond")
        return z

    def profile_iterations(self, b: torch.Tensor, precond: str = "ext_spai", iterations: int = 20):
        """Mean device ms per kernel class over `iterations` updates: [G^T r, G t (+rho), p, A p (+pq), x/r (+rr)]."""
        x = torch.empty_like(b)
        ms = (C.c_float * 5)()
        _lib.check(self._lib.b200spai_pcg_profile_iterations(self._h, PRECOND[precond], _ptr(b.contiguous()), _ptr(x),
                                                             int(iterations), ms, _stream()), "pcg_profile_iterations")
        return [float(v) for v in ms]

    def fused_profile(self):
        """Mode 4, after a solve: ([ms in t = G^T r, z = G t, p update, q = A p, x/r update] incl. the barrier closing
        each phase, iterations covered)."""
        ms = (C.c_double * 5)()
        it = C.c_int32()
        _lib.check(self._lib.b200spai_pcg_fused_profile(self._h, ms, C.byref(it), _stream()), "pcg_fused_profile")
        return [float(v) for v in ms], it.value

    @property
    def launches(self) -> int:
        v = C.c_int64()
        _lib.check(self._lib.b200spai_pcg_launch_count(self._h, C.byref(v)), "pcg_launch_count")
        return v.value

    def close(self):
        if self._h:
            self._lib.b200spai_pcg_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
