#!/bin/bash
# ncu launch list (per-kernel durations) of a short bench run: tools/ncu_list.sh TAG READS m1 m2 ...
TAG=$1; READS=$2; shift 2
mkdir -p gpurun_out
for m in "$@"; do
  SHORT="python bench.py --method $m --reads $READS --steps 1 --warmup 1 --methods= --no-e2e --no-cpu-baseline $EXTRA"
  $SHORT > gpurun_out/plain_${m}_$TAG.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches_${m}_$TAG.csv $SHORT > gpurun_out/ncu_list_${m}_$TAG.log 2>&1
  echo "$m ncu exit $?"
  python tools/launch_list.py gpurun_out/launches_${m}_$TAG.csv 2>&1 | tail -25
done
