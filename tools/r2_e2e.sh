#!/bin/bash
O=gpurun_out; T=${1:-r02q}
python bench.py --steps 6 --warmup 2 --methods "" --no-cpu-baseline --reads 60000 > $O/bench_e2e_$T.json 2> $O/bench_e2e_$T.err; echo "bench exit $?"; tail -3 $O/bench_e2e_$T.err
python -c "
import json; b=json.load(open('$O/bench_e2e_$T.json')); e=b['e2e']; print('e2e', e['value'], 'serial', e['serial_value'], e['serial_compress_gbs'], e['serial_decompress_gbs'], 'ceiling', e['link_ceiling_gbs'], e['link'])"
