#!/bin/bash
# 1-GPU reference points of the row-partitioned bench (216^3, 256^3) + the default bench
mkdir -p gpurun_out
for G in ${GRIDS:-216 256}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port $((29700+G)) bench.py --mode rowpart --gpus 1 --steps 3 --warmup 2 --grid $G 2> gpurun_out/rowpart_1_$G.err | grep '^{' | tee gpurun_out/rowpart_1_$G.json | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print(d['n_gpus'], 'ms/step', d['ms_per_step'], 'ms/iter', d['ms_per_iteration'], d['iterations'], d['phases_ms_per_step'])"
  tail -2 gpurun_out/rowpart_1_$G.err | cut -c1-300
done
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench_default.json
