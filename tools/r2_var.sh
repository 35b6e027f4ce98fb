#!/bin/bash
# variant libraries against the in-tree one, no verification (diagnostic builds may write garbage)
O=gpurun_out; M=${M:-vbe21_zd}
for v in "$@" base; do
  L=""; [ $v != base ] && L="PRESS_GPU_LIB=$PWD/variants/lib_$v.so"
  env $L timeout 300 python bench.py --method $M --reads 125000 --steps 3 --warmup 2 --methods "" --no-cpu-baseline --no-e2e --no-verify > $O/var_$v.json 2> $O/var_$v.err || { echo "$v FAILED"; tail -3 $O/var_$v.err; continue; }
  python -c "
import json; b=json.load(open('$O/var_$v.json')); print('$v', round(b['compress_gbs']), round(b['decompress_gbs']), {k:round(x['frac'],3) for k,x in b['kernels'].items()})"
done
