"""The inference benchmark driver (learningsparsepreconditioner4gpu_b200/infer.py, reference infer.py): table / CSV schema
and right-hand-side modes on CPU; the whole sample loop on the GPU."""
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from learningsparsepreconditioner4gpu_b200 import infer  # noqa: E402


def test_timestat_tables_match_reference_schema(tmp_path):
    st = infer.Timestat()
    st.put("PCG-none-cuda", 0.010, 0.0, 100, 512)
    st.put("PCG-none-cuda", 0.030, 0.0, 300, 1024)
    st.put("Neural+CUDA", 0.004, 0.002, 40, 512)
    df = st.timestat_to_dataframe()
    assert list(df.columns) == ["Key", "Total Time (ms)", "Solve Time (ms)", "Precond Time (ms)", "#Iteration"]   # infer.py:104-110
    row = df[df.Key == "PCG-none-cuda"].iloc[0]
    assert row["Solve Time (ms)"] == 20.0 and row["#Iteration"] == 200.0 and row["Total Time (ms)"] == 20.0
    row = df[df.Key == "Neural+CUDA"].iloc[0]
    assert row["Total Time (ms)"] == 6.0 and row["Precond Time (ms)"] == 2.0
    al = st.all_time_stat()
    assert list(al.columns) == ["Key", "Solve Time (ms)", "Precond Time (ms)", "#Iteration", "Matrix Size"]       # infer.py:135-141
    assert len(al) == 3 and al["Matrix Size"].tolist() == [512.0, 1024.0, 512.0]
    f1, f2 = infer.save_stats(st, tmp_path / "out", "heat", 1e-6, infer_prefix="x_")
    assert f1.name == "infer_x_heat_6.csv" and f2.name == "all_infer_x_heat_6.csv"                                # infer.py:374-381
    import pandas as pd

    back = pd.read_csv(f1)
    assert back.Key.tolist() == ["PCG-none-cuda", "Neural+CUDA"]
    st.print(out=lambda s: None)


def test_rhs_modes():
    mask = torch.tensor([[1.0], [0.0], [1.0], [1.0]])
    m = mask.numpy().reshape(-1)
    assert np.array_equal(infer.make_rhs("mask", mask), m) and np.array_equal(infer.make_rhs("ones", mask), m)
    r = infer.make_rhs("random", mask, rng=np.random.default_rng(0))
    assert r[1] == 0.0 and np.all(r[[0, 2, 3]] != 0.0)
    A_full = sp.csr_matrix(np.ones((4, 4)))
    nb = infer.make_rhs("neighbour", mask, A_full)
    assert np.allclose(nb, (A_full @ (1 - m) + 0.1 * m) * m) and nb[1] == 0.0                                      # infer.py:304-305
    with pytest.raises(ValueError):
        infer.make_rhs("neighbour", mask)
    with pytest.raises(ValueError):
        infer.make_rhs("nope", mask)


def test_get_workspace_names():
    from learningsparsepreconditioner4gpu_b200.neural_cg.utils import validate

    cls, fn = infer.get_workspace("simple")
    assert cls.__name__ == "SimpleTrainingWorkspace" and fn is validate.get_pcg_iter_time
    cls, fn = infer.get_workspace("scaled")
    assert cls.__name__ == "ScaledTrainingWorkspace" and fn is validate.get_pcg_scaled_iter_time
    with pytest.raises(ValueError):
        infer.get_workspace("other")                                                                               # infer.py:163
    with pytest.raises(ValueError):
        infer.run_inference(None, [], device="cpu")


def test_run_inference_fails_loudly_without_cuda():
    from learningsparsepreconditioner4gpu_b200 import _lib, systems

    if torch.cuda.is_available():
        pytest.skip("needs a CPU-only box")
    s = systems.poisson_sample(5, seed=1)
    ws = systems.default_workspace(s)
    with pytest.raises((_lib.B200SpaiError, RuntimeError, AssertionError)):      # .to("cuda") or require_cuda, never a CPU solve
        infer.run_inference(ws, [s], warmup=0)


@pytest.mark.gpu
@pytest.mark.parametrize("rhs", ["mask", "neighbour"])
def test_run_inference_on_gpu(tmp_path, rhs):
    from learningsparsepreconditioner4gpu_b200 import systems
    from oracle import assemble as oasm
    from oracle import pcg as opcg

    samples = [systems.poisson_sample(g, seed=5 + g) for g in (8, 10)]
    ws = systems.default_workspace(samples[0])
    stats = infer.run_inference(ws, samples, rhs=rhs, rtol=1e-6, repeat=2, warmup=2)
    assert set(stats.stat_dict) == {"PCG-none-cuda", "PCG-diagonal-cuda", "Neural+CUDA"}
    neural = stats.stat_dict["Neural+CUDA"]
    assert neural.all_matrix_size == [8 ** 3, 10 ** 3] and all(t > 0 for t in neural.all_prec_time)
    # the neural iteration counts are the oracle's (reference get_pcg_iter_time_scipy) to +-1
    for s, it in zip(samples, neural.all_iteration):
        n = s.num_nodes
        G, _ = ws.inference_step(s.to("cuda"))
        A = oasm.to_csr_cpu(s.edge_index, s.matrix_values, n, s.mask)
        gt = s.mask.numpy().reshape(-1).astype(np.float64)
        if rhs == "neighbour":
            A_full = oasm.to_csr_cpu(s.edge_index, s.matrix_values, n, None)
            A_full.data.fill(1.0)
            gt = (A_full @ (1 - gt) + 0.1 * gt) * gt
        assert abs(it - opcg.pcg_scipy(A, gt, G, ws.epsilon, rtol=1e-6)) <= 1
    assert all(i > 0 for i in stats.stat_dict["PCG-none-cuda"].all_iteration)
    f1, f2 = infer.save_stats(stats, tmp_path, "poisson3d", 1e-6)
    assert f1.exists() and f2.exists()
