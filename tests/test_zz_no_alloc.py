"""GPU test of the allocation-free preconditioner apply (b200spai_spai_apply_noalloc, csrc/noalloc.cu): an entry point
without a counterpart in the reference, added for callers that cannot call cudaMalloc.

Collected LAST on purpose (a failure here must not hide other tests) and independent of
test order: it builds its own matrices and passes alone in a fresh process as well as after the other tests.

The output is checked against the existing entry point (b200spai_pcg_apply_precond, "ext_spai") on two calls in a row with
different inputs -- the second call finds the first call's intermediate in the shared static scratch -- within 1e-12 relative
(the bar the residual-history tests use; both paths add a row's products in ascending column order)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL = 1e-12
MAX_ROWS = 1 << 20   # B200SPAI_NOALLOC_MAX_ROWS (include/b200spai.h)


def _apply_noalloc(lib, G, GT, eps, r, z):
    return lib.b200spai_spai_apply_noalloc(G._h, GT._h, float(eps), C.c_void_p(r.data_ptr()), C.c_void_p(z.data_ptr()),
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream))


@pytest.mark.parametrize("shape", ["poisson", "elast_b3"])
def test_noalloc_apply_matches_existing_entry_point(shape):
    from learningsparsepreconditioner4gpu_b200 import _lib, systems
    from learningsparsepreconditioner4gpu_b200.device import DevicePCG
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils.validate import to_csr_device

    lib = _lib.load()
    s, b = (systems.poisson_sample(20, seed=3, device="cuda"), 1) if shape == "poisson" else \
           (systems.elasticity_sample(12, 6, 6, seed=3, device="cuda"), 3)
    n, eps = s.num_nodes * b, 3.0e-3
    gen = torch.Generator(device="cuda").manual_seed(11)
    g_vals = 0.1 + 0.5 * torch.randn(s.num_edges, b, b, device="cuda", generator=gen)
    A = to_csr_device(s.edge_index, s.matrix_values, n, s.mask)
    G = to_csr_device(s.edge_index, g_vals, n, s.mask)
    GT = G.transpose()
    solver = DevicePCG(A)
    solver.set_spai(G, eps)
    rs = [torch.randn(n, dtype=torch.float64, device="cuda", generator=gen) for _ in range(2)]
    refs = [solver.apply_precond(r, "ext_spai") for r in rs]
    zs = [torch.full((n,), float("nan"), dtype=torch.float64, device="cuda") for _ in rs]
    # set-up call: the first launch from csrc/noalloc.cu makes the driver load that module (CUDA loads modules lazily), which
    # is when the static scratch array comes to exist on this GPU -- the 8 MiB the header states, not an allocation per call
    _lib.check(_apply_noalloc(lib, G, GT, eps, rs[0], zs[0]), "spai_apply_noalloc")
    zs[0].fill_(float("nan"))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    cached0 = C.c_int64()
    _lib.check(lib.b200spai_pool_cached_bytes(C.byref(cached0)), "pool_cached_bytes")
    for r, z in zip(rs, zs):                     # two calls in a row, nothing in between
        _lib.check(_apply_noalloc(lib, G, GT, eps, r, z), "spai_apply_noalloc")
    torch.cuda.synchronize()
    cached1 = C.c_int64()
    _lib.check(lib.b200spai_pool_cached_bytes(C.byref(cached1)), "pool_cached_bytes")
    assert torch.cuda.mem_get_info()[0] == free0 and cached1.value == cached0.value   # no device memory changed hands
    for k, (z, ref) in enumerate(zip(zs, refs)):
        err = float((z - ref).abs().max() / ref.abs().max())
        print(f"{shape}: call {k}: n={n} max|dz|/max|z| = {err:.2e}")
        assert torch.isfinite(z).all() and err <= RTOL, (shape, k, err)
    assert not torch.equal(zs[0], zs[1])
    # and against scipy's own arithmetic on the host
    Gs = G.to_scipy()
    ref_host = Gs @ (Gs.T.tocsr() @ rs[1].cpu().numpy()) + eps * rs[1].cpu().numpy()
    assert np.abs(zs[1].cpu().numpy() - ref_host).max() <= 1e-12 * np.abs(ref_host).max()


def test_noalloc_apply_rejects_sizes_beyond_the_static_scratch():
    from learningsparsepreconditioner4gpu_b200 import _lib
    from learningsparsepreconditioner4gpu_b200.device import DeviceCSR

    lib = _lib.load()
    n = MAX_ROWS + 1
    ip = torch.arange(n + 1, dtype=torch.int32, device="cuda")
    big = DeviceCSR.from_torch(ip, ip[:-1].clone(), torch.ones(n, dtype=torch.float64, device="cuda"), (n, n))
    r = torch.ones(n, dtype=torch.float64, device="cuda")
    z = torch.zeros_like(r)
    rc = _apply_noalloc(lib, big, big, 1.0, r, z)
    assert rc == -2                                              # B200SPAI_ERR_ARG
    msg = lib.b200spai_last_error().decode()
    assert "B200SPAI_NOALLOC_MAX_ROWS" in msg and str(n) in msg, msg
    torch.cuda.synchronize()
    assert float(z.abs().max()) == 0.0                           # nothing was launched
    # exactly at the limit it runs: identity matrix -> z = (1 + eps) r
    ip = ip[:MAX_ROWS + 1].clone()
    eye = DeviceCSR.from_torch(ip, ip[:-1].clone(), torch.ones(MAX_ROWS, dtype=torch.float64, device="cuda"), (MAX_ROWS, MAX_ROWS))
    r, z = r[:MAX_ROWS].clone(), z[:MAX_ROWS].clone()
    _lib.check(_apply_noalloc(lib, eye, eye, 0.5, r, z), "spai_apply_noalloc")
    torch.cuda.synchronize()
    assert torch.equal(z, 1.5 * r)
