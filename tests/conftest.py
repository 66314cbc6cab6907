import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["poisson_b1", "poisson_zeros_b1", "heat_b1", "elast_b3", "synthetic_b1", "synthetic_mild_b1"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    import numpy as np

    return np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False)


def golden_state_dict(g):
    import torch

    return {str(k): torch.from_numpy(g["sd/" + str(k)].copy()) for k in g["sd_keys"]}


def golden_csr(g, prefix):
    from scipy.sparse import csr_matrix

    n = len(g[prefix + "_indptr"]) - 1
    return csr_matrix((g[prefix + "_data"], g[prefix + "_indices"], g[prefix + "_indptr"]), shape=(n, n))


TRAIN_CASES = ["poisson_b1", "heat_b1", "elast_b3", "synthetic_mild_b1", "poisson_pair_b1"]


def load_train_case(name):
    """tests/golden/train_<name>.npz (minted from the unmodified reference by make_golden_train.py) as
    (batch: dict of CPU tensors incl. residual / ptr / batch, state_dict, fixture)."""
    import numpy as np
    import torch

    t = load_golden("train_" + name)
    gs = [load_golden(str(c)) for c in t["cases"]]
    offs = np.cumsum([0] + [g["x"].shape[0] for g in gs])

    def cat(k):
        return torch.cat([torch.from_numpy(g[k]) for g in gs])

    batch = dict(x=cat("x"), edge_attr=cat("edge_attr"), mask=cat("mask"), matrix_values=cat("matrix_values"),
                 inv_diag=cat("inv_diag"), rsqrt_diag=cat("rsqrt_diag"),
                 edge_index=torch.cat([torch.from_numpy(g["raw_edge_index"]) + int(o) for g, o in zip(gs, offs[:-1])], dim=1),
                 residual=torch.from_numpy(t["residual"]), ptr=torch.from_numpy(t["ptr"]),
                 batch=torch.repeat_interleave(torch.arange(len(gs)), torch.as_tensor(np.diff(offs))))
    sd = {str(k): torch.from_numpy(t["sd/" + str(k)].copy()) for k in t["sd_keys"]}
    return batch, sd, t


@pytest.fixture(params=GOLDEN_CASES)
def golden(request):
    return request.param, load_golden(request.param)
