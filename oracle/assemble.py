"""ORACLE (test infrastructure, not product): CPU restatement of the reference's CSR assembly.

Follows neural_cg/utils/validate.py:13-51 (to_numpy, to_csr_cpu), neural_cg/data.py:134-156
(make_bsr_from_coo_inds) and neural_cg/data.py:159-170 (apply_dbc_masking), written out explicitly
instead of through scipy's sparse add so that the pattern rules are visible:

* scalar path: COO -> CSR, duplicates summed, stored zeros KEPT when ``mask is None``;
* block path: the b*b blocks are attached to the *sorted* block-CSR structure in the order given
  (the reference assumes row-major / column-ascending input, data.py:150-156) and expanded to scalars;
* masking (``mask is not None``): every entry whose row or column DOF has mask == 0 is zeroed, a unit
  diagonal is added on those DOFs, and -- because the reference does this with a scipy sparse add --
  every stored entry that is exactly 0.0 afterwards disappears from the pattern;
* result: fp64 CSR with int32 indices sorted within rows.

Parity status: PINNED -- tests/test_oracle_golden.py compares indptr/indices/data bit-for-bit against
the reference's own to_csr_cpu run in the build container (tests/golden/make_golden.py).
"""
from __future__ import annotations

import numpy as np
import torch
from scipy.sparse import csr_matrix


def to_numpy(x) -> np.ndarray:
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    if isinstance(x, np.ndarray):
        return x
    raise ValueError(f"Unknown type {type(x)}")


def expand_blocks(row, col, vals):
    """[E,b,b] blocks at (row,col) -> scalar COO triplets, row-major inside each block."""
    b = vals.shape[-1]
    ii, jj = np.meshgrid(np.arange(b), np.arange(b), indexing="ij")
    r = (row[:, None, None] * b + ii[None]).reshape(-1)
    c = (col[:, None, None] * b + jj[None]).reshape(-1)
    return r, c, vals.reshape(-1)


def to_csr_cpu(edge_index, edge_attr, n: int, mask, dtype=np.float64) -> csr_matrix:
    row = to_numpy(edge_index[0]).astype(np.int32)
    col = to_numpy(edge_index[1]).astype(np.int32)
    vals = to_numpy(edge_attr).astype(dtype)
    assert vals.ndim in (1, 3)
    if vals.ndim == 3 and vals.shape[1] > 1:
        b = vals.shape[-1]
        nb = n // b
        # block structure is the sorted CSR of the block graph; blocks are attached in input order
        structure = csr_matrix((np.ones(row.size), (row, col)), shape=(nb, nb))
        brow = np.repeat(np.arange(nb, dtype=np.int64), np.diff(structure.indptr))
        r, c, v = expand_blocks(brow, structure.indices.astype(np.int64), vals)
        mat = csr_matrix((v, (r, c)), shape=(n, n), dtype=dtype)
    else:
        mat = csr_matrix((vals.reshape(-1), (row, col)), shape=(n, n), dtype=dtype)
    if mask is not None:
        m = to_numpy(mask).reshape(-1).astype(dtype)
        coo = mat.tocoo()
        keep = (m[coo.row] != 0) & (m[coo.col] != 0)
        data = np.where(keep, coo.data, 0.0)
        fixed = np.nonzero(m == 0)[0]
        ident = 1.0 - m  # reference adds diags(1 - mask): exactly 1 on masked DOFs when mask is 0/1
        r = np.concatenate([coo.row, np.arange(n)])
        c = np.concatenate([coo.col, np.arange(n)])
        d = np.concatenate([data, ident])
        mat = csr_matrix((d, (r, c)), shape=(n, n), dtype=dtype)  # duplicates summed
        mat.eliminate_zeros()
        del fixed
    mat = csr_matrix(mat)
    mat.sort_indices()
    return mat


def to_csr_cpu_sorted(edge_index, edge_attr, n: int, mask, dtype=np.float64) -> csr_matrix:
    """Same matrix as ``to_csr_cpu`` for the scalar case (b = 1) when the edge list is what the reference's
    datasets produce (row-major, ascending columns, duplicate-free, neural_cg/data.py:442-450), built directly
    from the sorted triplets in O(E) numpy passes -- for the >= 10 M-row CPU baseline in bench.py, where the
    COO round trips of the general function cost minutes.  Falls back to ``to_csr_cpu`` when a masked row has no
    stored diagonal.  Equality with ``to_csr_cpu`` is tested on the fixtures (tests/test_oracle_golden.py)."""
    row = to_numpy(edge_index[0]).astype(np.int64)
    col = to_numpy(edge_index[1]).astype(np.int64)
    vals = to_numpy(edge_attr)
    assert vals.size == row.size, "scalar systems only"
    data = vals.reshape(-1).astype(dtype)
    key = row * n + col
    if row.size and not (np.diff(key) > 0).all():
        return to_csr_cpu(edge_index, edge_attr, n, mask, dtype)
    indptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(row, minlength=n), out=indptr[1:])
    if mask is not None:
        m = to_numpy(mask).reshape(-1).astype(dtype)
        on_diag = row == col
        has_diag = np.zeros(n, dtype=bool)
        has_diag[row[on_diag]] = True
        if not has_diag[m != 1].all():
            return to_csr_cpu(edge_index, edge_attr, n, mask, dtype)
        data = np.where((m[row] != 0) & (m[col] != 0), data, 0.0)
        data[on_diag] += (1.0 - m)[row[on_diag]]
    mat = csr_matrix((data, col.astype(np.int32), indptr.astype(np.int32)), shape=(n, n))
    if mask is not None:
        mat.eliminate_zeros()
    mat.has_sorted_indices = True
    return mat
