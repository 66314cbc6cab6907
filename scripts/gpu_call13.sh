#!/bin/bash
# 8-GPU: partitioned-system worker at world 8 (bit-exactness vs single GPU), row-partitioned bench at 216^3 and 256^3
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29650 tests/dist_gnn_worker.py poisson 32 2>&1 | grep '^{' | cut -c1-400
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29651 tests/dist_worker.py 32 2 2>&1 | grep '^{' | cut -c1-600
export B200SPAI_DIST_PROFILE=1
for G in ${GRIDS:-216 256}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700+G)) bench.py --mode rowpart --gpus $N --steps 3 --warmup 2 --grid $G 2> gpurun_out/rowpart_${N}_$G.err | grep '^{' | tee gpurun_out/rowpart_${N}_$G.json | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print(d['n_gpus'], 'ms/step', d['ms_per_step'], 'ms/iter', d['ms_per_iteration'], d['iterations'], d['phases_ms_per_step'])"
  grep "per-launch" gpurun_out/rowpart_${N}_$G.err | head -3
  tail -2 gpurun_out/rowpart_${N}_$G.err | cut -c1-300
done
