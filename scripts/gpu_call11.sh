#!/bin/bash
# row-partitioned bench with the per-launch profile of the distributed iteration
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
export B200SPAI_DIST_PROFILE=1
for W in ${WS:-1 2}; do
  if [ $W -le $N ]; then
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port $((29700+W)) bench.py --mode rowpart --gpus $W --steps 2 --warmup 2 --grid ${GRID:-160} 2> gpurun_out/rowpart_$W.err | grep '^{' | tee gpurun_out/rowpart_$W.json | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print(d['n_gpus'], 'ms/step', d['ms_per_step'], 'ms/iter', d['ms_per_iteration'], d['iterations'], d['phases_ms_per_step'])"
    grep "per-launch" gpurun_out/rowpart_$W.err
  fi
done
