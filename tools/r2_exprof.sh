#!/bin/bash
# launch list + full captures of the exception encoders, digested on the box
TAG=$1; M=${2:-vbe21_zd}; shift 2
bash tools/ncu_list.sh $TAG 20000 $M | grep -E "ex_|ncu exit"
bash tools/ncu_full.sh $TAG 20000 $M "$@" | grep "ncu exit"
for K in "$@"; do
  python -c "
import sys; sys.path.insert(0, 'tools'); import profile_digest as pd
pd.kernels('gpurun_out/full_${M}_${K}_$TAG.ncu-rep', 'gpurun_out/digest_${M}_${K}_$TAG.txt')"
  head -60 gpurun_out/digest_${M}_${K}_$TAG.txt | cut -c1-150
  ncu -i gpurun_out/full_${M}_${K}_$TAG.ncu-rep --page source --csv > gpurun_out/sass_${M}_${K}_$TAG.csv 2>/dev/null
done
rm -f gpurun_out/*.ncu-rep
