"""neural_cg/dataset.py (FolderDataset, FolderWriter, to_bcoo_components) against fixtures produced by the UNMODIFIED
reference reader (tests/golden/make_golden_dataset.py): same folders, field-by-field equality.  CPU only."""
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from learningsparsepreconditioner4gpu_b200.neural_cg.dataset import FolderDataset, FolderWriter, to_bcoo_components  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
FIELDS = ("x", "edge_index", "edge_attr", "matrix_values", "mask", "diagonal", "inv_diag", "rsqrt_diag")
CASES = {
    "heat_fixed": dict(is_fixed_topology=True, block_size=1, has_shared_features=True, use_node_features=True,
                       use_matrix_as_edge_feature=True, use_mask_as_node_feature=True, use_node_features_as_edge_feature=False,
                       use_edge_features_as_node_feature="disable", use_random_rhs=False, normalize_matrix="mean"),
    "poisson_free": dict(is_fixed_topology=False, block_size=1, has_shared_features=False, use_node_features=True,
                         use_matrix_as_edge_feature=True, use_mask_as_node_feature=True, use_node_features_as_edge_feature=False,
                         use_edge_features_as_node_feature="mean", use_random_rhs=True, normalize_matrix="mean"),
    "elast_free": dict(is_fixed_topology=False, block_size=3, has_shared_features=False, use_node_features=True,
                       use_matrix_as_edge_feature=True, use_mask_as_node_feature=True, use_node_features_as_edge_feature=False,
                       use_edge_features_as_node_feature="disable", use_random_rhs=True, normalize_matrix="mean"),
}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "dataset_folders.npz"))


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("in_memory", [False, True])
def test_reader_matches_reference(gold, name, in_memory):
    ds = FolderDataset(load_into_memory=in_memory, prefix=os.path.join(GOLD, "folders", name), **CASES[name])
    assert len(ds) == int(gold[f"{name}/len"])
    assert ds.num_node_features == int(gold[f"{name}/num_node_features"])
    assert ds.num_edge_features == int(gold[f"{name}/num_edge_features"])
    for i in range(len(ds)):
        d = ds[i]
        for f in FIELDS:
            key = f"{name}/{i}/{f}"
            if key not in gold:
                continue
            ours, ref = getattr(d, f).numpy(), gold[key]
            assert ours.shape == ref.shape and ours.dtype == ref.dtype, (key, ours.shape, ref.shape, ours.dtype, ref.dtype)
            assert np.array_equal(ours, ref), key
        # the edge list is what the device path needs: row-major, columns ascending, no duplicates
        ei = d.edge_index.numpy()
        k = ei[0].astype(np.int64) * (ei.max() + 1) + ei[1]
        assert np.all(np.diff(k) > 0)


def test_bcoo_components_first_appearance_order_and_overwrite():
    # scalar row 1 touches a block column that row 0 does not: the reference appends that block later
    rows = np.array([0, 0, 1, 1, 1, 2, 3, 3])
    cols = np.array([0, 4, 1, 2, 5, 2, 3, 3])
    vals = np.arange(1.0, 9.0)
    coo = sp.coo_matrix((vals, (rows, cols)), shape=(4, 6))
    bv, br, bc = to_bcoo_components(coo, 2)
    assert list(zip(br, bc)) == [(0, 0), (0, 2), (0, 1), (1, 1)]
    assert bv[0].tolist() == [[1.0, 0.0], [0.0, 3.0]] and bv[2].tolist() == [[0.0, 0.0], [4.0, 0.0]]
    assert bv[3].tolist() == [[6.0, 0.0], [0.0, 8.0]]          # duplicate (3,3): the last value wins
    with pytest.raises(TypeError):
        to_bcoo_components(coo.tocsr(), 2)
    with pytest.raises(ValueError):
        to_bcoo_components(coo, 0)


def test_writer_reader_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    n = 30
    M = sp.random(n, n, density=0.15, random_state=1, format="csr")
    A = sp.csr_matrix(M @ M.T + sp.identity(n))
    A.sort_indices()
    w = FolderWriter(str(tmp_path / "ds"), block_size=1, is_fixed_topology=True, save_rhs=3, save_lhs=True)
    w.write_topology(A)
    mask = (rng.random((n, 1)) > 0.2).astype(np.float64)
    w.append(A, mask=mask, features=rng.standard_normal((n, 2)), seed=1)
    w.append(2.0 * A, mask=mask, features=rng.standard_normal((n, 2)), seed=2)
    ds = FolderDataset(is_fixed_topology=True, load_into_memory=False, block_size=1, has_shared_features=False,
                       use_node_features=True, use_matrix_as_edge_feature=True, use_mask_as_node_feature=True,
                       use_node_features_as_edge_feature=False, use_edge_features_as_node_feature="disable",
                       use_random_rhs=False, normalize_matrix=False, prefix=str(tmp_path / "ds"))
    assert len(ds) == 6 and ds.samples[4] == (1, 1)
    raw = ds.get_internal(4)
    assert np.array_equal(raw.block_values.reshape(-1), 2.0 * A.data)
    assert np.allclose((2.0 * A) @ raw.lhs.reshape(-1), raw.rhs.reshape(-1), atol=1e-10)
    assert np.array_equal(raw.rhs.reshape(-1) * mask.reshape(-1), raw.rhs.reshape(-1))   # rhs lives on the free DOFs
    d = ds[0]
    assert d.x.shape == (n, 3) and d.edge_attr.shape == (A.nnz, 1)


def test_reader_errors(tmp_path):
    kw = dict(is_fixed_topology=False, load_into_memory=False, block_size=1, has_shared_features=False,
              use_node_features=False, use_matrix_as_edge_feature=True, use_mask_as_node_feature=True,
              use_node_features_as_edge_feature=False, use_edge_features_as_node_feature="disable", use_random_rhs=True,
              normalize_matrix="mean")
    (tmp_path / "mat").mkdir()
    with pytest.raises(FileNotFoundError):
        FolderDataset(prefix=str(tmp_path), **kw)
    with pytest.raises(ValueError):
        FolderDataset(prefix=os.path.join(GOLD, "folders", "poisson_free"),
                      **{**kw, "use_node_features_as_edge_feature": True, "use_edge_features_as_node_feature": "mean"})
    with pytest.raises(FileNotFoundError):      # fixed topology without demo.mtx
        FolderDataset(prefix=os.path.join(GOLD, "folders", "poisson_free"), **{**kw, "is_fixed_topology": True})
