#!/bin/bash
# parity tests + SpMV configuration sweep (delta16 on/off, stream configs) at 216^3
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/pytest_gpu.log
: > gpurun_out/sweep.jsonl
for cfg in 0 1 2 3 4; do
  B200SPAI_STREAM_CFG=$cfg timeout 300 python scripts/spmv_sweep.py --grid 216 >> gpurun_out/sweep.jsonl 2>> gpurun_out/sweep.err
done
B200SPAI_NO_COL16=1 timeout 300 python scripts/spmv_sweep.py --grid 216 >> gpurun_out/sweep.jsonl 2>> gpurun_out/sweep.err
cat gpurun_out/sweep.jsonl
tail -5 gpurun_out/sweep.err
