#!/bin/bash
# ncu: launch list of the GNN forward + full capture of the edge kernel (grid 128^3)
mkdir -p gpurun_out
CMD="python bench.py --grid 128 --steps 1 --warmup 1 --no-cpu-baseline"
export B200SPAI_PCG_MODE=1
timeout 600 $CMD > gpurun_out/plain128.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gnn_ -c 40 --csv --log-file gpurun_out/launches_gnn128.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gnn_edge -s 5 -c 2 -o gpurun_out/prof_gnn_edge $CMD > gpurun_out/ncu_gnn.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_gnn.log
