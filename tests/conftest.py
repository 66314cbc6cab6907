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


@pytest.fixture(params=GOLDEN_CASES)
def golden(request):
    return request.param, load_golden(request.param)
