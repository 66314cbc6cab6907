#!/bin/bash
# GNN code-size change: parity tests, infer-flow tests, gnn phase at 160^3 and 216^3
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gnn_gpu.py tests/test_infer_flow.py -m gpu -q 2>&1 | tail -4
for G in 160 216; do
  timeout 300 python bench.py --grid $G --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['iterations'], d['phases_ms_per_step'])"
done
