#!/bin/bash
# Is the second sweep of the exception encoders worth keeping in L2?  ncu launch lists of a batch
# that fits L2 (len-cap 8192 x READS reads), caches flushed between kernels (ncu's default) and not.
TAG=$1; READS=${2:-3000}; CAP=${3:-8192}; M=${4:-vbe21_zd}
mkdir -p gpurun_out
SHORT="python bench.py --method $M --reads $READS --len-cap $CAP --steps 1 --warmup 1 --methods= --no-e2e --no-cpu-baseline"
$SHORT > gpurun_out/l2probe_plain_$TAG.log 2>&1 || { tail -5 gpurun_out/l2probe_plain_$TAG.log; exit 1; }
tail -1 gpurun_out/l2probe_plain_$TAG.log | cut -c1-400
for cc in all none; do
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control $cc -c 400 --csv \
      --log-file gpurun_out/l2probe_${cc}_$TAG.csv $SHORT > gpurun_out/l2probe_ncu_${cc}_$TAG.log 2>&1
  echo "cache-control $cc: ncu exit $?"
  grep -E "ex_count|ex_emit|ex_decode2|ex_size|ex_header|ex_offsets" gpurun_out/l2probe_${cc}_$TAG.csv | awk -F'","' '{print $5, $(NF-2), $(NF-1), $NF}' | tr -d '"' | head -40
done
