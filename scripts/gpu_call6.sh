#!/bin/bash
# multi-GPU: dist tests, row-partitioned bench at the visible GPU count, batch bench at the same count
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N"
[ -n "$SKIP_TESTS" ] || timeout 600 python -m pytest tests/test_dist_gpu.py -m gpu -q -x 2>&1 | tail -4
for W in ${WS:-1 2 4 8}; do
  if [ $W -le $N ]; then
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port $((29700+W)) bench.py --mode rowpart --gpus $W --steps 2 --warmup 1 --grid ${GRID:-216} 2> gpurun_out/rowpart_$W.err | grep '^{' | tee gpurun_out/rowpart_$W.json | cut -c1-1600
    tail -2 gpurun_out/rowpart_$W.err
  fi
done
