#!/bin/bash
O=gpurun_out; T=${1:-r02g}
for cap in 0 300000; do
for m in svb16_zd shuff_vbe21_zd vbe21_zd; do
python bench.py --method $m --reads 125000 --steps 3 --warmup 2 --methods "" --no-cpu-baseline --no-e2e --len-cap $cap > $O/cap_${m}_${cap}_$T.json 2> $O/cap_${m}_${cap}_$T.err
python -c "
import json; b=json.load(open('$O/cap_${m}_${cap}_$T.json')); print('$m cap $cap', round(b['compress_gbs']), round(b['decompress_gbs']), {k:round(v['frac'],3) for k,v in b['kernels'].items()})"
done; done
