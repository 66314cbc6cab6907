"""How repetitive are the rows of the C5 matrices?  (CPU analysis behind DESIGN section 8, "row-pattern dictionary")

    python scripts/row_pattern_stats.py [grid=96]

For the assembled A (= the pattern of G; G^T likewise) of the jittered Kuhn-tet Poisson system: the within-row column steps the
16-bit stream stores (first entry: 0) form one tuple per row.  Reports how many distinct tuples there are, how many rows the
most frequent ones cover, and what the column stream would cost with a dictionary (one u16 pattern id per row; rows whose
pattern is not among the first 65 535 keep their explicit steps) against the 2 bytes per entry of today."""
import json
import os
import sys
from collections import Counter

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from learningsparsepreconditioner4gpu_b200 import systems
    from oracle import assemble as oasm

    grid = int(sys.argv[1]) if len(sys.argv) > 1 else 96
    s = systems.poisson_sample(grid, seed=100, gen_device="cpu", device="cpu")
    n = s.num_nodes
    A = oasm.to_csr_cpu_sorted(s.edge_index, s.matrix_values, n, s.mask)
    out = {"grid": grid, "n": n}
    for name, M in (("A", A), ("A^T (pattern of G^T)", A.T.tocsr())):
        M.sort_indices()
        ip, ix = M.indptr.astype(np.int64), M.indices.astype(np.int64)
        steps = np.diff(ix, prepend=0)
        steps[ip[:-1][np.diff(ip) > 0]] = 0                      # first entry of every row
        lens = np.diff(ip)
        # hash every row's step tuple: pad to the longest row, view rows as bytes
        L = int(lens.max())
        pad = np.zeros((n, L), dtype=np.int32)
        rows = np.repeat(np.arange(n), lens)
        pos = np.arange(ix.size) - np.repeat(ip[:-1], lens)
        pad[rows, pos] = steps
        pad = np.concatenate([lens[:, None].astype(np.int32), pad], axis=1)
        keys = np.ascontiguousarray(pad).view(np.dtype((np.void, pad.dtype.itemsize * pad.shape[1]))).ravel()
        uniq, counts = np.unique(keys, return_counts=True)
        order = np.argsort(-counts)
        cum = np.cumsum(counts[order])
        nnz = int(ix.size)
        top = {str(k): round(float(cum[min(k, cum.size) - 1]) / n, 4) for k in (1, 4, 16, 64, 256, 65535)}
        covered_rows = int(cum[min(65535, cum.size) - 1])
        covered_entries = int(lens[np.isin(keys, uniq[order[:65535]])].sum())
        dict_bytes = 2 * n + 2 * (nnz - covered_entries) + int((lens[np.isin(keys, uniq[order[:65535]], invert=False)] * 0).sum())
        dict_table = int(sum(2 * (len(bytes(u)) // 4 - 1) for u in uniq[order[:65535]][:1]) * min(65535, uniq.size))
        out[name] = {"nnz": nnz, "mean_row": round(nnz / n, 2), "distinct_patterns": int(uniq.size), "rows_covered_by_top_k": top,
                     "rows_in_dictionary": round(covered_rows / n, 4),
                     "col_stream_bytes_per_entry_now": 2.0,
                     "col_stream_bytes_per_entry_dictionary": round((dict_bytes + dict_table) / nnz, 3)}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
