#!/bin/bash
# The one parametrised GPU-box job (run through gpurun; everything it writes goes to gpurun_out/).
#   scripts/gpu_job.sh <tag> <step> [<step> ...]
# steps:  tests[:<pytest args>]   pytest -m gpu (default: the whole suite)
#         bench[:<bench args>]    python bench.py <args> (1 GPU) -> gpurun_out/<tag>_bench*.json
#         tbench:<N>[:<args>]     torchrun with N ranks: bench.py --gpus N <args>
#         launches[:<args>]       ncu launch list of bench.py <args> (gpu__time_duration, no clock control)
#         ncu:<kernel regex>[:<args>]  one ncu --set full capture of the first launches matching the regex
#         py:<script and args>    python <script>
set -u
mkdir -p gpurun_out
TAG=$1; shift
PORT=29700
for step in "$@"; do
  kind=${step%%:*}; rest=""; [[ "$step" == *:* ]] && rest=${step#*:}
  echo "=== [$TAG] $step"
  case $kind in
    tests)
      timeout 2400 python -m pytest ${rest:-tests} -m gpu -x -q -s 2>&1 | tail -40 | tee gpurun_out/${TAG}_pytest.log ;;
    bench)
      timeout 1500 python bench.py $rest 2> gpurun_out/${TAG}_bench.err | grep '^{' | tee -a gpurun_out/${TAG}_bench.json | cut -c1-1500
      tail -5 gpurun_out/${TAG}_bench.err ;;
    tbench)
      n=${rest%%:*}; a=""; [[ "$rest" == *:* ]] && a=${rest#*:}
      PORT=$((PORT+1))
      timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT \
        bench.py --gpus $n $a 2> gpurun_out/${TAG}_bench_${n}.err | grep '^{' | tee -a gpurun_out/${TAG}_bench_${n}.json | cut -c1-1500
      tail -5 gpurun_out/${TAG}_bench_${n}.err ;;
    launches)
      timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${TAG}_launches.csv \
        python bench.py $rest > gpurun_out/${TAG}_launches.log 2>&1
      python scripts/summarize_launches.py gpurun_out/${TAG}_launches.csv | tee gpurun_out/${TAG}_launches.txt | head -40 ;;
    ncu)
      k=${rest%%:*}; a=""; [[ "$rest" == *:* ]] && a=${rest#*:}
      timeout 1500 ncu --set full --clock-control none --import-source on -k "regex:$k" -s 2 -c 3 -o gpurun_out/${TAG}_ncu -f \
        python bench.py $a > gpurun_out/${TAG}_ncu.log 2>&1
      ncu -i gpurun_out/${TAG}_ncu.ncu-rep --page raw --csv > gpurun_out/${TAG}_ncu_raw.csv 2>/dev/null
      ls -la gpurun_out/${TAG}_ncu* ;;
    py)
      timeout 2400 python $rest 2>&1 | tail -60 | tee gpurun_out/${TAG}_py.log ;;
    *) echo "unknown step $step" ;;
  esac
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader | head -8
