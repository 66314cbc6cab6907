"""ctypes binding of libb200spai.so (the C ABI declared in include/b200spai.h).

The library is REQUIRED: there is no CPU or eager-PyTorch fallback behind these calls.  If it is
missing it is built with nvcc (``build.py``); if that fails, importing raises.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200SPAI_LIB", os.path.join(HERE, "libb200spai.so"))

c_i32, c_i64, c_f64, c_vp = C.c_int32, C.c_int64, C.c_double, C.c_void_p


class PcgResult(C.Structure):
    _fields_ = [
        ("iterations", c_i32), ("converged", c_i32), ("breakdown", c_i32), ("reserved", c_i32),
        ("b_norm", c_f64), ("final_residual", c_f64), ("solve_ms", C.c_float), ("setup_ms", C.c_float),
    ]


# name -> (argtypes); every function returns int except last_error
SIGNATURES = {
    "b200spai_version": [],
    "b200spai_device_info": [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)],
    "b200spai_total_launches": [C.POINTER(c_i64)],
    "b200spai_csr_create_host": [c_i32, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_csr_create_dev": [c_i32, c_i32, c_i64, c_vp, c_vp, c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_csr_transpose": [c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_csr_shape": [c_vp, C.POINTER(c_i32), C.POINTER(c_i32), C.POINTER(c_i64)],
    "b200spai_csr_download": [c_vp, c_vp, c_vp, c_vp, c_vp],
    "b200spai_csr_device_ptrs": [c_vp, C.POINTER(c_vp), C.POINTER(c_vp), C.POINTER(c_vp)],
    "b200spai_csr_destroy": [c_vp],
    "b200spai_spmv": [c_vp, c_vp, c_vp, C.c_int, c_vp],
    "b200spai_csr_diagonal": [c_vp, c_vp, c_vp],
    "b200spai_edges_to_csr": [c_i64, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_pcg_create": [c_vp, C.POINTER(c_vp)],
    "b200spai_pcg_set_spai": [c_vp, c_vp, c_f64, c_vp],
    "b200spai_pcg_set_mode": [c_vp, C.c_int],
    "b200spai_pcg_solve_dev": [c_vp, C.c_int, c_vp, c_vp, c_f64, c_i32, c_vp, c_vp, C.POINTER(PcgResult)],
    "b200spai_pcg_solve_host": [c_vp, C.c_int, c_vp, c_vp, c_f64, c_i32, c_vp, c_vp, C.POINTER(PcgResult)],
    "b200spai_pcg_apply_precond": [c_vp, C.c_int, c_vp, c_vp, c_vp],
    "b200spai_pcg_profile_iterations": [c_vp, C.c_int, c_vp, c_vp, c_i32, C.POINTER(C.c_float), c_vp],
    "b200spai_pcg_launch_count": [c_vp, C.POINTER(c_i64)],
    "b200spai_pcg_fused_profile": [c_vp, C.POINTER(c_f64), C.POINTER(c_i32), c_vp],
    "b200spai_pcg_refresh": [c_vp, c_vp],
    "b200spai_csr_bsr_info": [c_vp, c_vp, C.POINTER(c_i32), C.POINTER(c_i64)],
    "b200spai_edges_to_csr_plan": [c_i64, c_i64, c_i32, c_vp, c_vp, c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_csr_has_plan": [c_vp, C.POINTER(c_i32)],
    "b200spai_csr_refresh_from_edges": [c_vp, c_vp, c_i64, c_vp, C.POINTER(c_i32)],
    "b200spai_csr_transpose_plan": [c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_csr_refresh_transposed": [c_vp, c_vp, c_vp],
    "b200spai_pcg_destroy": [c_vp],
    "b200spai_graph_create": [c_i64, c_i64, c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_graph_destroy": [c_vp],
    "b200spai_gnn_create": [c_i32, c_i32, c_i32, c_i32, c_vp, c_i64, C.POINTER(c_vp)],
    "b200spai_gnn_destroy": [c_vp],
    "b200spai_gnn_forward": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp],
    "b200spai_gnn_forward_stage": [c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp],
    "b200spai_gnn_node_features": [c_vp, c_i32, C.POINTER(c_vp)],
    "b200spai_gnn_launch_count": [c_vp, C.POINTER(c_i64)],
    "b200spai_graph_spmv_f32": [c_vp, c_i32, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp],
    "b200spai_gnn_tape_create": [c_vp, c_vp, C.POINTER(c_vp)],
    "b200spai_gnn_tape_destroy": [c_vp],
    "b200spai_gnn_forward_train": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp],
    "b200spai_gnn_backward": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp],
    "b200spai_gnn_set_weights_dev": [c_vp, c_vp, c_i64, c_vp],
    "b200spai_graph_outer_f32": [c_vp, c_i32, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp],
    "b200spai_csr_stream_info": [c_vp, C.POINTER(c_i32), C.POINTER(c_i32), c_vp],
    "b200spai_spai_apply_noalloc": [c_vp, c_vp, c_f64, c_vp, c_vp, c_vp],
    "b200spai_pool_trim": [],
    "b200spai_pool_cached_bytes": [C.POINTER(c_i64)],
    "b200spai_dist_create": [c_i32, c_i32, c_i64, C.POINTER(c_vp)],
    "b200spai_dist_ipc_handle": [c_vp, c_vp],
    "b200spai_dist_connect": [c_vp, c_vp, c_vp],
    "b200spai_dist_relayout": [c_vp, c_vp],
    "b200spai_dist_info": [c_vp, C.POINTER(c_i32), C.POINTER(c_i32), C.POINTER(c_i64), C.POINTER(c_i32)],
    "b200spai_csr_slice_rows": [c_vp, c_i32, c_i32, c_i64, c_i32, c_vp, C.POINTER(c_vp)],
    "b200spai_dist_pcg_setup": [c_vp, c_vp, c_vp, c_vp, c_f64],
    "b200spai_dist_pcg_set_mode": [c_vp, C.c_int],
    "b200spai_dist_pcg_fused_ok": [c_vp, c_vp, C.POINTER(c_i32)],
    "b200spai_dist_pcg_fused_profile": [c_vp, C.POINTER(c_f64), C.POINTER(c_i32), c_vp],
    "b200spai_dist_pcg_fused_barrier_profile": [c_vp, C.POINTER(c_f64), C.POINTER(c_i64), c_vp],
    "b200spai_dist_pcg_solve": [c_vp, C.c_int, c_vp, c_vp, c_f64, c_i32, c_vp, c_vp, C.POINTER(PcgResult)],
    "b200spai_dist_pcg_profile_iterations": [c_vp, C.c_int, c_vp, c_vp, c_i32, C.POINTER(C.c_float), c_vp],
    "b200spai_dist_destroy": [c_vp],
    "b200spai_pcg_solve_batch": [c_vp, C.c_int, c_i32, C.POINTER(c_i32), c_vp, c_vp, c_f64, c_i32, c_vp, c_i64, c_vp,
                                 C.POINTER(PcgResult)],
}

_lib = None


class B200SpaiError(RuntimeError):
    pass


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load (building first if needed) libb200spai.so and attach the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise B200SpaiError(f"{LIB_PATH} is missing; run python -m learningsparsepreconditioner4gpu_b200.build")
        from . import build as _build

        _build.build(verbose=False)
    lib = C.CDLL(LIB_PATH)
    lib.b200spai_last_error.restype = C.c_char_p
    lib.b200spai_last_error.argtypes = []
    missing = [name for name in SIGNATURES if not hasattr(lib, name)]
    if missing:
        raise B200SpaiError(f"{LIB_PATH} does not export {missing}; rebuild it (python -m learningsparsepreconditioner4gpu_b200.build --force)")
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = C.c_int
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().b200spai_last_error().decode("utf-8", "replace")
        raise B200SpaiError(f"{what} failed with status {rc}: {msg}")


def total_launches() -> int:
    """Kernel launches made by libb200spai in this process so far."""
    v = c_i64()
    check(load().b200spai_total_launches(C.byref(v)), "total_launches")
    return v.value


def require_cuda():
    """Fail loudly when there is no CUDA device: nothing here has a CPU path."""
    import torch

    if not torch.cuda.is_available():
        raise B200SpaiError("libb200spai needs a CUDA device (sm_100a); no CPU fallback exists")


class nvtx_range:
    """NVTX range around a phase of the path (``with nvtx_range("b200spai/gnn"): ...``) for nsys / ncu timelines.  Off unless
    B200SPAI_NVTX=1 (a push / pop pair costs ~1 us of host time: nothing inside a timed step by default)."""

    enabled = os.environ.get("B200SPAI_NVTX", "0") not in ("", "0")

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if nvtx_range.enabled:
            import torch

            torch.cuda.nvtx.range_push(self.name)
        return self

    def __exit__(self, *exc):
        if nvtx_range.enabled:
            import torch

            torch.cuda.nvtx.range_pop()
        return False
