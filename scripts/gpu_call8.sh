#!/bin/bash
# ncu full capture of the GNN message-passing kernel (grid 128^3)
mkdir -p gpurun_out
CMD="python bench.py --grid 128 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain128.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gnn_mp -s 4 -c 2 -o gpurun_out/prof_gnn_mp $CMD > gpurun_out/ncu_gnn.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_gnn.log
ls -la gpurun_out
