#!/bin/bash
# GNN kernel variants: parity tests on the default, then the gnn phase at 160^3 for each B200SPAI_GNN_CFG
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gnn_gpu.py -m gpu -q 2>&1 | tail -4
for cfg in ${CFGS:-0 2 4 6}; do
  echo "== cfg $cfg"
  B200SPAI_GNN_CFG=$cfg timeout 300 python bench.py --grid ${GRID:-160} --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['phases_ms_per_step'])"
done
