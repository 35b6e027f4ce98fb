#!/bin/bash
# one `ncu --set full` capture per kernel regex: tools/ncu_full.sh TAG READS METHOD regex1 regex2 ...
TAG=$1; READS=$2; M=$3; shift 3
mkdir -p gpurun_out
SHORT="python bench.py --method $M --reads $READS --steps 1 --warmup 1 --methods= --no-e2e --no-cpu-baseline $EXTRA"
$SHORT > gpurun_out/plain_full_${M}_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_full_${M}_$TAG.log; exit 1; }
for K in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 1 -c 1 \
      -f -o gpurun_out/full_${M}_${K}_$TAG $SHORT > gpurun_out/ncu_full_${M}_${K}_$TAG.log 2>&1
  echo "$K ncu exit $?"
done
ls -la gpurun_out/*.ncu-rep
