#!/bin/bash
# Refresh of one evidence round for the exception methods only (after a change to their kernels):
# bench lines, the ex-zd sweep, launch list and captures of the vbe21 kernels.  tools/gpu_round_ex.sh TAG
TAG=${1:-r02c}; O=gpurun_out; mkdir -p $O
for m in vbe21_zd vbsse21_zd vbbe21_zd ex_zd; do
  python bench.py --method $m --methods "" --reads 125000 --steps 3 --warmup 3 --e2e-gb 1.0 --cpu-reads 1500 --cpu-seconds 4 \
      > $O/bench_${m}_$TAG.json 2> $O/bench_${m}_$TAG.err; echo "$m exit $?"
done
bash tools/sweep.sh $TAG 20000 > $O/sweep_$TAG.log 2>&1; tail -6 $O/sweep_$TAG.log
bash tools/r2_exprof.sh $TAG vbe21_zd ex_count ex_emit_bytes ex_decode2 | grep -E "ncu exit|time_duration"
