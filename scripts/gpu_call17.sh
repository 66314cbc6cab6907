#!/bin/bash
# round-end style validation + refreshed launch list: all GPU tests, smoke, default bench, reference arm, ncu launch list (stream mode)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_default.json')); print({k:d[k] for k in ['value','ms_per_step','iterations','phases_ms_per_step','gpu_launches','clocks']}); print(d['e2e']); print({k:d['roofline'][k] for k in ['achieved','frac','streamed','ms_per_launch']}); print(d['cpu_baseline'])"
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
timeout 300 python bench.py --mode batch_small --steps 3 --warmup 2 > gpurun_out/bench_batch_small.json 2>/dev/null; cut -c100-400 gpurun_out/bench_batch_small.json
export B200SPAI_PCG_MODE=1
CMD="python bench.py --grid 216 --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s 30000 -c 12000 --csv --log-file gpurun_out/launches_grid216.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
