#!/bin/bash
# 2-GPU: all GPU tests (incl. partitioned-system / multi-rank) + row-partitioned bench at 1 and 2 ranks
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l); echo "GPUs: $N"
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -15
for W in ${WS:-1 2}; do
  if [ $W -le $N ]; then
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port $((29700+W)) bench.py --mode rowpart --gpus $W --steps 3 --warmup 2 --grid ${GRID:-160} 2> gpurun_out/rowpart_$W.err | grep '^{' | tee gpurun_out/rowpart_$W.json | cut -c1-1800
    tail -3 gpurun_out/rowpart_$W.err
  fi
done
