#!/bin/bash
# 8-GPU closing run: row-partitioned 216^3 (final kernels), C2 batch of small systems, default bench
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
export B200SPAI_DIST_PROFILE=1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29916 bench.py --mode rowpart --gpus $N --steps 3 --warmup 2 --grid 216 2> gpurun_out/rowpart_${N}_216.err | grep '^{' | tee gpurun_out/rowpart_${N}_216.json | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print(d['n_gpus'], 'ms/step', d['ms_per_step'], 'ms/iter', d['ms_per_iteration'], d['iterations'], d['phases_ms_per_step'])"
grep "per-launch" gpurun_out/rowpart_${N}_216.err | head -1
unset B200SPAI_DIST_PROFILE
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29917 bench.py --mode batch_small --gpus $N --batch 512 --steps 3 --warmup 2 2>/dev/null | grep '^{' | tee gpurun_out/bench_batch_small_$N.json | cut -c100-420
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29918 bench.py --gpus $N --steps 3 --warmup 3 2>/dev/null | grep '^{' | tee gpurun_out/bench_default_$N.json | cut -c100-330
