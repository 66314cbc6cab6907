"""The kernel variants `auto` does not pick on the fixtures stay under the same bars: both small-system PCG kernels forced
on every golden system (pcg_small.cuh / pcg_small2.cuh), and the all-FFMA GNN kernels next to the tensor-core default."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _run(what, **env):
    e = dict(os.environ)
    e.update({k: str(v) for k, v in env.items()})
    r = subprocess.run([sys.executable, os.path.join(HERE, "alt_config_worker.py"), what], env=e, capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0 and "ALT-CONFIG-OK" in r.stdout, (r.stdout[-2000:], r.stderr[-4000:])


@pytest.mark.parametrize("kernel", [1, 2])
def test_small_system_kernels_forced(kernel):
    _run("small", B200SPAI_SMALL_KERNEL=kernel)


def test_ffma_gnn_kernels():
    _run("gnn", B200SPAI_GNN_CFG=2)
