"""Dataset wire format of the reference: the ``generated/<exp_name>/`` folders written by
``DatagenBase`` (neural_cg/datagen_helper.py:199-322) and read by ``FolderDataset`` (neural_cg/data.py:339-640).

    <prefix>/mat/000000.npy | .mtx      fixed topology: CSR value vector of the scalar matrix (pattern in demo.mtx);
                                        otherwise one MatrixMarket file per system            (datagen_helper.py:254-260)
    <prefix>/demo.mtx                   the shared sparsity pattern (fixed topology)          (datagen_helper.py:349-352)
    <prefix>/mask/000000.npy            [N, b]  1 = free DOF, 0 = Dirichlet                   (:269-273)
    <prefix>/features/000000.npy        [N, k]  per-sample node features                      (:264-268)
    <prefix>/shared_features.npy        [N, k'] node features shared by every sample          (:353-355)
    <prefix>/rhs/000000.npy, lhs/...    [n, m]  m right-hand sides / solutions per matrix     (:303-321)

``FolderDataset`` keeps the reference's constructor keywords (``config/data.yaml`` splats in unchanged) and returns the
``Data`` objects of ``neural_cg.data.make_data``; ``FolderWriter`` is the writer side so that systems produced here
(``datagen.py``) and folders produced by the reference are interchangeable.  Host code only: files, scipy, numpy.
Edge lists come out row-major / column-ascending -- the order ``b200spai_edges_to_csr`` and the GNN require and the
order scipy's ``coo_matrix(csr)`` produces for the reference (data.py:442-450,500-503).
"""
from __future__ import annotations

from pathlib import Path
from typing import List, Optional, Sequence, Tuple, Union

import numpy as np
from scipy.io import mmread, mmwrite
from scipy.sparse import coo_matrix, csr_matrix

from .data import Data, RawData, make_data


def to_bcoo_components(coo: coo_matrix, block_size: int) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Scalar COO -> (block_values [NB, b, b], block_row [NB], block_col [NB]) (data.py:15-66).  Blocks appear in order
    of the first stored entry that touches them, like the reference's dict accumulation; later duplicates overwrite."""
    if not isinstance(coo, coo_matrix):
        raise TypeError("Input must be a scipy.sparse.coo_matrix")
    if block_size <= 0:
        raise ValueError("Block size must be positive")
    b = int(block_size)
    br, bc = coo.row // b, coo.col // b
    ncb = int(coo.shape[1] + b - 1) // b
    key = br.astype(np.int64) * ncb + bc.astype(np.int64)
    uniq, first, inv = np.unique(key, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")            # first-appearance order of the blocks
    rank = np.empty_like(order)
    rank[order] = np.arange(order.size)
    slot = rank[inv]
    vals = np.zeros((uniq.size, b, b))
    vals[slot, coo.row % b, coo.col % b] = coo.data      # numpy assigns in index order: the last duplicate wins
    return vals, (uniq[order] // ncb).astype(int), (uniq[order] % ncb).astype(int)


def _mmread(path) -> csr_matrix:
    try:
        return csr_matrix(mmread(path, spmatrix=True))
    except TypeError:  # scipy < 1.15 has no spmatrix keyword
        return csr_matrix(mmread(path))


def _sorted(files) -> List[Path]:
    out = list(files)
    out.sort()
    return out


class FolderDataset:
    """Reader of one dataset folder; same keywords and sample semantics as the reference's ``FolderDataset``."""

    def __init__(self, is_fixed_topology: bool, load_into_memory: bool, block_size: int, has_shared_features: bool,
                 use_node_features: bool, use_matrix_as_edge_feature: bool, use_mask_as_node_feature: bool,
                 use_node_features_as_edge_feature: bool, use_edge_features_as_node_feature: str, use_random_rhs: bool,
                 normalize_matrix: Union[bool, str], prefix: str):
        self.is_fixed_topology, self.block_size, self.prefix = bool(is_fixed_topology), int(block_size), prefix
        self.path = Path(prefix)
        mat_dir = self.path / "mat"
        self.all_matrices = _sorted(mat_dir.glob("*.npy")) if self.is_fixed_topology else _sorted(mat_dir.glob("*.mtx"))
        self.all_lhs = _sorted((self.path / "lhs").glob("*.npy"))
        self.all_rhs = _sorted((self.path / "rhs").glob("*.npy"))
        self.all_masks = _sorted(self.path.glob("mask/*.npy"))
        self.all_features = _sorted((self.path / "features").glob("*.npy"))
        if not self.all_matrices:
            raise FileNotFoundError(f"no matrices under {mat_dir}")
        for name, lst in (("lhs", self.all_lhs), ("rhs", self.all_rhs)):
            if lst and len(lst) != len(self.all_matrices):
                raise ValueError(f"{len(lst)} {name} files for {len(self.all_matrices)} matrices")
        self.has_shared_features = bool(has_shared_features)
        self.shared_features = np.load(self.path / "shared_features.npy") if self.has_shared_features else None
        # one sample per (matrix, right-hand-side column)  (data.py:393-398)
        self.samples: List[Tuple[int, int]] = []
        for idx, f in enumerate(self.all_rhs):
            cols = np.load(f, mmap_mode="r")
            self.samples += [(idx, i) for i in range(cols.shape[1] if cols.ndim > 1 else 1)]
        self.use_node_features = bool(use_node_features)
        self.use_matrix_as_edge_feature = bool(use_matrix_as_edge_feature)
        self.use_mask_as_node_feature = bool(use_mask_as_node_feature)
        self.use_node_features_as_edge_feature = bool(use_node_features_as_edge_feature)
        self.use_edge_features_as_node_feature = use_edge_features_as_node_feature
        self.use_random_rhs, self.normalize_matrix = bool(use_random_rhs), normalize_matrix
        if self.use_node_features_as_edge_feature and self.use_edge_features_as_node_feature != "disable":
            raise ValueError("You cannot enable both feature enhancers")
        # feature widths (data.py:413-433)
        nf = 0
        if self.use_node_features:
            if len(self.all_features) != len(self.all_matrices):
                raise ValueError("use_node_features needs one features/*.npy per matrix")
            nf = np.load(self.all_features[0], mmap_mode="r").shape[1]
            if self.shared_features is not None:
                nf += self.shared_features.shape[1]
        if self.use_mask_as_node_feature:
            nf += self.block_size
        ef = self.block_size ** 2 if self.use_matrix_as_edge_feature else 0
        if self.use_node_features_as_edge_feature:
            ef += 2 * nf
        if self.use_edge_features_as_node_feature != "disable":
            nf += ef
        self.num_node_features_, self.num_edge_features_ = nf, ef
        if self.is_fixed_topology:
            topo = self.path / "demo.mtx"
            if not topo.exists():
                raise FileNotFoundError(f"fixed topology needs {topo}")
            self.topo_mat_dofs = _mmread(topo).sorted_indices()
            self.edge_index = self._block_pattern(self.topo_mat_dofs)
        self.loaded: List[RawData] = []
        if load_into_memory:
            self.loaded = [self.get_internal(i) for i in range(len(self))]

    # ---- sizes --------------------------------------------------------------------------------------------
    @property
    def num_node_features(self) -> int:
        return self.num_node_features_

    @property
    def num_edge_features(self) -> int:
        return self.num_edge_features_

    def len(self) -> int:
        return max(len(self.all_matrices), len(self.samples))

    __len__ = len

    # ---- loading ------------------------------------------------------------------------------------------
    def _block_pattern(self, m: csr_matrix) -> np.ndarray:
        """[2, NB] block (= graph) pattern of a scalar matrix, row-major / column-ascending (data.py:438-450)."""
        b = self.block_size
        bsr = m.tobsr((b, b))
        bsr.sort_indices()
        rows = np.repeat(np.arange(bsr.shape[0] // b), np.diff(bsr.indptr))
        return np.vstack((rows, bsr.indices)).astype(np.int64)

    def load(self, mat_file: Path, lhs_file: Optional[Path], rhs_file: Optional[Path], feature_file: Optional[Path],
             mask_file: Optional[Path], dtype=np.float64) -> RawData:
        b = self.block_size
        if str(mat_file).endswith(".npy"):                       # values only; pattern from demo.mtx (data.py:474-486)
            if not self.is_fixed_topology:
                raise ValueError(".npy matrices need is_fixed_topology")
            topo = self.topo_mat_dofs
            matrix = csr_matrix((np.load(mat_file), topo.indices, topo.indptr), shape=topo.shape)
            bsr = matrix.tobsr((b, b))
            bsr.sort_indices()
            block_values, edge_index = bsr.data.astype(dtype).copy(), self.edge_index
            num_nodes = topo.shape[0] // b
        else:                                                    # a whole matrix per sample (data.py:487-503)
            matrix = _mmread(mat_file)
            coo = coo_matrix(matrix)
            if b == 1:
                block_values = coo.data.astype(dtype).reshape(-1, 1, 1).copy()
                edge_index = np.vstack((coo.row, coo.col)).astype(np.int64)
            else:
                block_values, br, bc = to_bcoo_components(coo, b)
                edge_index = np.vstack((br, bc)).astype(np.int64)
            num_nodes = matrix.shape[0] // b
        diagonals = matrix.diagonal().reshape(-1, b)
        lhs = rhs = node_features = None
        if rhs_file is not None:
            rhs = np.load(rhs_file)
            if rhs.ndim == 1:
                rhs = rhs.reshape(-1, 1)
            elif rhs.ndim > 2:
                raise ValueError(f"Unexpected RHS shape: {rhs.shape}")
            if lhs_file is not None:
                lhs = np.load(lhs_file)
                if lhs.ndim == 1:
                    lhs = lhs.reshape(-1, 1)
                elif lhs.ndim > 2:
                    raise ValueError(f"Unexpected LHS shape: {lhs.shape}")
                assert lhs.shape == rhs.shape
        if self.use_node_features:
            parts = []
            if feature_file is not None:
                feat = np.load(feature_file)
                assert feat.ndim == 2 and feat.shape[0] == num_nodes
                parts.append(feat)
            if self.shared_features is not None:
                parts.append(self.shared_features)
            node_features = np.concatenate(parts, axis=-1)
        mask = np.ones((num_nodes, b), dtype=dtype)
        if mask_file is not None:
            loaded = np.load(mask_file)
            assert loaded.shape == mask.shape, (loaded.shape, mask.shape)
            mask = loaded
        return RawData(block_values, diagonals, edge_index, node_features, lhs, rhs, mask, num_nodes, b)

    def get_internal(self, idx: int) -> RawData:
        if self.loaded:
            return self.loaded[idx]
        mat_id, sub_id = self.samples[idx] if self.samples else (idx, 0)
        pick = lambda lst: lst[mat_id] if lst else None  # noqa: E731
        raw = self.load(self.all_matrices[mat_id], pick(self.all_lhs), pick(self.all_rhs), pick(self.all_features),
                        pick(self.all_masks))
        b = self.block_size
        sub = lambda v: v[:, sub_id].reshape(-1, b) if v is not None else None  # noqa: E731
        return RawData(raw.block_values, raw.diagonals, raw.edge_index, raw.node_features, sub(raw.lhs), sub(raw.rhs),
                       raw.mask, raw.num_nodes, raw.block_size)

    def get(self, idx: int) -> Data:
        data = make_data(self.get_internal(idx), self.use_matrix_as_edge_feature, self.use_mask_as_node_feature,
                         self.use_node_features_as_edge_feature, self.use_edge_features_as_node_feature,
                         self.use_random_rhs, self.normalize_matrix, is_inference=False)
        assert data.x.shape[-1] == self.num_node_features and data.edge_attr.shape[-1] == self.num_edge_features
        return data

    __getitem__ = get

    def __iter__(self):
        return (self.get(i) for i in range(len(self)))


class FolderWriter:
    """Writer side of the format (``DatagenBase.prepare/append/generate``, datagen_helper.py:231-322,346-355)."""

    def __init__(self, prefix: str, block_size: int = 1, is_fixed_topology: bool = False, save_rhs: int = 0,
                 save_lhs: bool = False):
        self.path, self.block_size, self.is_fixed_topology = Path(prefix), int(block_size), bool(is_fixed_topology)
        self.save_rhs, self.save_lhs, self.current_count = int(save_rhs), bool(save_lhs), 0
        for sub in ("mask", "mat", "features") + (("rhs",) if self.save_rhs else ()) + (("lhs",) if self.save_lhs else ()):
            (self.path / sub).mkdir(parents=True, exist_ok=True)

    def write_topology(self, topo: csr_matrix) -> None:
        mmwrite(self.path / "demo.mtx", csr_matrix(topo).sorted_indices())

    def write_shared_features(self, feats: np.ndarray) -> None:
        np.save(self.path / "shared_features.npy", np.asarray(feats))

    def append(self, mat: csr_matrix, mask: Optional[np.ndarray] = None, features: Optional[np.ndarray] = None,
               rhs: Union[np.ndarray, Sequence[np.ndarray], None] = None, seed: Optional[int] = None) -> csr_matrix:
        mat = csr_matrix(mat)
        rows, name = mat.shape[0], f"{self.current_count:06d}"
        if self.is_fixed_topology:
            np.save(self.path / "mat" / f"{name}.npy", mat.sorted_indices().data)
        else:
            mmwrite(self.path / "mat" / f"{name}.mtx", mat)
        if features is not None:
            features = features.reshape(-1, 1) if features.ndim == 1 else features
            assert features.shape[0] == rows // self.block_size
            np.save(self.path / "features" / f"{name}.npy", features)
        if mask is not None:
            mask = mask.reshape(-1, 1) if mask.ndim == 1 else mask
            assert mask.shape[0] == rows // self.block_size
            np.save(self.path / "mask" / f"{name}.npy", mask)
        if self.save_rhs:
            if isinstance(rhs, np.ndarray):
                cols = [rhs.reshape(-1)]
            elif rhs is not None:
                cols = [np.asarray(v).reshape(-1) for v in rhs]
            else:                                             # random unit right-hand sides on the free DOFs (:286-292)
                rng = np.random.default_rng(seed)
                cols = []
                for _ in range(self.save_rhs):
                    v = rng.standard_normal(rows)
                    v /= np.linalg.norm(v)
                    cols.append(v * mask.reshape(-1) if mask is not None else v)
            stacked = np.stack(cols, axis=0).astype(np.float64)
            np.save(self.path / "rhs" / f"{name}.npy", stacked.T)
            if self.save_lhs:
                from scipy.sparse.linalg import splu

                lu = splu(mat.tocsc())
                np.save(self.path / "lhs" / f"{name}.npy", np.stack([lu.solve(v) for v in stacked], axis=0).T)
        self.current_count += 1
        return mat
