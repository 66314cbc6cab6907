"""Pin the CPU oracle against the golden fixtures minted from the reference's own code
(tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

from conftest import golden_csr, golden_state_dict
from oracle import assemble, data as odata, gnn as ognn, pcg as opcg


def _csr_equal(a, b):
    return (np.array_equal(a.indptr, b.indptr) and np.array_equal(a.indices, b.indices)
            and np.array_equal(a.data, b.data))


def test_make_data_matches_reference(golden):
    name, g = golden
    b = int(g["block_size"])
    nf = g["raw_node_features"]
    d = odata.make_data(
        g["raw_values"], g["raw_edge_index"], g["raw_mask"], None if nf.shape[1] == 0 else nf, g["raw_diag"],
        use_edge_features_as_node_feature=str(g["feat_edge_as_node"]),
    )
    for k in ("x", "edge_attr", "matrix_values", "mask", "inv_diag", "rsqrt_diag", "diagonal"):
        assert np.array_equal(d[k].numpy(), g[k]), (name, k)
    assert d["matrix_values"].shape[1:] == (b, b)


def test_gnn_forward_matches_reference(golden):
    name, g = golden
    sd = golden_state_dict(g)
    b = int(g["block_size"])
    G = ognn.workspace_forward(sd, torch.from_numpy(g["x"]), torch.from_numpy(g["raw_edge_index"]),
                               torch.from_numpy(g["edge_attr"]), b)
    # same torch CPU kernels in the same order on both sides -> bit-identical
    assert np.array_equal(G.numpy(), g["G"]), name
    assert set(ognn.state_dict_keys()) == set(sd.keys())


def test_message_norm_scale_is_inert(golden):
    name, g = golden
    sd = golden_state_dict(g)
    for k in list(sd):
        if k.endswith("node_msg_norm.scale"):
            sd[k] = sd[k] * 123.0
    G = ognn.workspace_forward(sd, torch.from_numpy(g["x"]), torch.from_numpy(g["raw_edge_index"]),
                               torch.from_numpy(g["edge_attr"]), int(g["block_size"]))
    assert np.array_equal(G.numpy(), g["G"]), name


def test_aatpe_matches_reference(golden):
    name, g = golden
    ei = torch.from_numpy(g["raw_edge_index"])
    G = torch.from_numpy(g["G"])
    x = torch.from_numpy(g["aatpe_in"])
    mask = torch.from_numpy(g["mask"])
    y = ognn.aatpe(x, ei, G, float(g["epsilon"]), mask)
    assert np.array_equal(y.numpy(), g["aatpe_out"]), name
    yd = ognn.aatpe(x, ei, G, float(g["epsilon"]), mask, torch.from_numpy(g["inv_diag"]))
    assert np.array_equal(yd.numpy(), g["aatpe_diag_out"]), name


def test_to_csr_matches_reference_bit_exact(golden):
    name, g = golden
    ei = torch.from_numpy(g["raw_edge_index"])
    n = g["mask"].size
    A = assemble.to_csr_cpu(ei, torch.from_numpy(g["matrix_values"]), n, torch.from_numpy(g["mask"]))
    assert _csr_equal(A, golden_csr(g, "A")), name
    A0 = assemble.to_csr_cpu(ei, torch.from_numpy(g["matrix_values"]), n, None)
    assert _csr_equal(A0, golden_csr(g, "A_nomask")), name
    G = assemble.to_csr_cpu(ei, torch.from_numpy(g["G"]), n, torch.from_numpy(g["mask"]))
    assert _csr_equal(G, golden_csr(g, "G")), name
    assert A.indices.dtype == np.int32 and A.data.dtype == np.float64


def test_sorted_fast_assembly_equals_general(golden):
    """to_csr_cpu_sorted (bench.py's CPU baseline at 10 M rows) is the same matrix as to_csr_cpu, bit for bit."""
    name, g = golden
    if int(g["block_size"]) != 1:
        pytest.skip("scalar systems only")
    ei = torch.from_numpy(g["raw_edge_index"])
    n = g["mask"].size
    for vals, key in ((g["matrix_values"], "A"), (g["G"], "G")):
        M = assemble.to_csr_cpu_sorted(ei, torch.from_numpy(vals), n, torch.from_numpy(g["mask"]))
        assert _csr_equal(M, golden_csr(g, key)), (name, key)
    M0 = assemble.to_csr_cpu_sorted(ei, torch.from_numpy(g["matrix_values"]), n, None)
    assert _csr_equal(M0, golden_csr(g, "A_nomask")), name


def test_csr_invariants(golden):
    """masked rows/cols are identity, stored zeros dropped, indices sorted."""
    name, g = golden
    A = golden_csr(g, "A")
    m = g["mask"].reshape(-1)
    assert A.has_sorted_indices and (A.data != 0).all()
    fixed = np.nonzero(m == 0)[0]
    D = A[fixed].tocoo()
    assert np.array_equal(np.sort(D.row), np.arange(len(fixed))) and (D.col == fixed[D.row]).all() and (D.data == 1).all()
    assert abs(A[:, fixed]).sum() == len(fixed)


def test_aatpe_consistent_with_csr_operator(golden):
    """AATPE(x) == G_csr (G_csr^T x) + eps x on free DOFs (fp32 vs fp64 tolerance)."""
    name, g = golden
    G = golden_csr(g, "G")
    m = g["mask"].reshape(-1).astype(np.float64)
    x = g["aatpe_in"].reshape(-1).astype(np.float64) * m
    ei = torch.from_numpy(g["raw_edge_index"])
    y32 = ognn.aatpe(torch.from_numpy((x.reshape(g["aatpe_in"].shape)).astype(np.float32)), ei,
                     torch.from_numpy(g["G"]), float(g["epsilon"]), torch.from_numpy(g["mask"])).numpy().reshape(-1)
    y64 = G @ (G.T @ x) + float(g["epsilon"]) * x
    free = m != 0
    assert np.allclose(y32[free], y64[free], rtol=2e-4, atol=2e-6), name


def test_pcg_restatement_matches_scipy_and_reference_counts(golden):
    name, g = golden
    A, G = golden_csr(g, "A"), golden_csr(g, "G")
    eps = float(g["epsilon"])
    gt = g["mask"].reshape(-1).astype(np.float64)
    b = A @ gt
    for kind, key in (("ext_spai", "iters_pcg"), ("none", "iters_cg"), ("ext_spai_scaled", "iters_pcg_scaled")):
        M = opcg.make_preconditioner(kind, A, G, eps)
        x, its, hist = opcg.pcg_restated(A, b, M, rtol=1e-6)
        assert its == int(g[key]), (name, kind, its, int(g[key]))
        assert opcg.pcg_scipy(A, gt, G, eps, kind=kind) == int(g[key])
        assert len(hist) == its + 1
    # history of the default solve: recurrence residual tracks the recorded true residual
    M = opcg.make_preconditioner("ext_spai", A, G, eps)
    x, its, hist = opcg.pcg_restated(A, b, M, rtol=1e-6)
    true_hist = g["true_res_hist"]
    assert len(true_hist) == its
    k = min(its, 25)
    assert np.allclose(hist[1 : k + 1], true_hist[:k], rtol=1e-6, atol=1e-12 * hist[0])
    if its < A.shape[0]:
        assert np.allclose(x, g["x_solution"], rtol=1e-9, atol=1e-12)
    _, its10, _ = opcg.pcg_restated(A, b, M, rtol=1e-10)
    assert its10 == int(g["iters_pcg_1e10"])


def test_pcg_zero_rhs():
    from scipy.sparse import identity

    x, its, hist = opcg.pcg_restated(identity(5, format="csr"), np.zeros(5), lambda r: r)
    assert its == 0 and hist == [0.0]


# ---- training step: oracle/train.py against the reference's own workspaces + loss under autograd --------
from conftest import TRAIN_CASES, load_train_case  # noqa: E402


@pytest.mark.parametrize("name", TRAIN_CASES)
@pytest.mark.parametrize("scaled", [False, True], ids=["simple", "scaled"])
def test_oracle_training_step_matches_reference(name, scaled):
    """Loss equal to the last bit, every parameter gradient within 2e-6 of max|grad| (autograd accumulation order)."""
    from oracle import train as otrain

    batch, sd, t = load_train_case(name)
    tag = "scaled" if scaled else "simple"
    loss, grads, _ = otrain.training_step(sd, batch, float(t["epsilon"]), int(t["block_size"]), scaled=scaled)
    assert float(loss) == float(np.float32(t[f"{tag}/loss"]))
    assert len(grads) == 114
    for k, g in grads.items():
        ref = t[f"{tag}/grad/{k}"]
        assert np.abs(g.numpy() - ref).max() <= 2e-6 * max(np.abs(ref).max(), 1e-12), k


LOSS_NAMES = ["RelativeL2Loss_PlainNorm", "RelativeL2Loss_ANorm", "L2Loss_ANorm", "PropLoss", "L1Loss", "RelPropLoss",
              "CosineSimilarityLoss_PlainNorm", "CosineSimilarityLoss_ANorm", "ConjGradLoss_PlainNorm", "ConjGradLoss_ANorm",
              "ConjGradLoss_ANorm_NoRelative", "NifLoss_Norm"]


@pytest.mark.parametrize("batch_less", [False, True], ids=["per_sample", "batch_less"])
@pytest.mark.parametrize("name", LOSS_NAMES)
def test_oracle_losses_match_reference(name, batch_less):
    """Every loss create_loss_item accepts: value and dL/dd against the reference's own classes (train_losses.npz)."""
    from conftest import load_golden
    from oracle import train as otrain

    batch, _, _ = load_train_case("poisson_pair_b1")
    g = load_golden("train_losses")
    batch["gt"] = torch.from_numpy(g["gt"])
    d = torch.from_numpy(g["d"]).clone().requires_grad_(True)
    val = otrain.loss_value(name, batch, d, [int(v) for v in batch["ptr"]], batch_less=batch_less)
    val.backward()
    tag = f"{name}/{'batch_less' if batch_less else 'per_sample'}"
    assert abs(float(val.detach()) - float(g[tag + "/loss"])) <= 2e-6 * abs(float(g[tag + "/loss"]))
    ref = g[tag + "/grad_d"]
    assert np.abs(d.grad.numpy() - ref).max() <= 2e-6 * np.abs(ref).max()
