"""Which driver / preconditioner kind deviates from the oracle's residual history on the synthetic SPD pair of
tests/test_c5_gpu.py::test_stream_epilogues_and_fallback_streams_vs_scipy (diagnostic, GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_c5_gpu import _spd_pair
from learningsparsepreconditioner4gpu_b200 import device as d
from oracle import pcg as opcg

for fp32_values, narrow in [(True, True), (False, False)]:
    A, G = _spd_pair(150000, 4, 7, fp32_values, narrow)
    Ad, Gd = d.DeviceCSR.from_scipy(A), d.DeviceCSR.from_scipy(G)
    rng = np.random.default_rng(1)
    rng.standard_normal(150000)
    b = A @ rng.standard_normal(150000)
    bd = torch.from_numpy(b).cuda()
    for kind in ("ext_spai", "ext_spai_scaled"):
        xo, its, ohist = opcg.pcg_restated(A, b, opcg.make_preconditioner(kind, A, G, 3e-3), rtol=1e-8)
        oh = np.asarray(ohist)
        for mode in (1, 2, 4):
            for rep in range(2):
                s2 = d.DevicePCG(Ad, mode=mode)
                s2.set_spai(Gd, 3e-3)
                x, res, hist = s2.solve(bd, kind, rtol=1e-8, history=True)
                h = hist.cpu().numpy()
                m = min(len(h), len(oh), 10)
                print(fp32_values, narrow, kind, "mode", mode, "rep", rep, "it", res.iterations, its,
                      "rel", np.array2string(np.abs(h[:m] - oh[:m]) / oh[:m], precision=1), flush=True)
