"""Runs golden-fixture GPU checks in a fresh process whose environment selects a NON-default kernel configuration (the
library reads these switches once per process):

    B200SPAI_SMALL_KERNEL=1|2 python tests/alt_config_worker.py small     # thread-per-row / lanes-per-row small-system PCG
    B200SPAI_GNN_CFG=2        python tests/alt_config_worker.py gnn       # all-FFMA GNN kernels

Prints ALT-CONFIG-OK on success.  Spawned by tests/test_alt_configs_gpu.py."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))


def main():
    what = sys.argv[1]
    from conftest import GOLDEN_CASES

    if what == "small":
        import test_pcg_gpu as t

        for name in GOLDEN_CASES:
            t.test_pcg_matches_reference_iterations(name, 3)
    elif what == "gnn":
        import test_gnn_gpu as t

        for name in GOLDEN_CASES:
            t.test_gnn_forward_matches_reference(name)
        t.test_gnn_forward_matches_oracle_larger(24, 1)
        t.test_gnn_forward_matches_oracle_larger(12, 3)
    else:
        raise SystemExit(f"unknown check {what}")
    print("ALT-CONFIG-OK")


if __name__ == "__main__":
    main()
