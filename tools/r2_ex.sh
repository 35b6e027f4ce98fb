#!/bin/bash
# exception-family check: GPU parity tests, then the 125k-read legs
O=gpurun_out; TAG=${1:-x}; shift
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu_$TAG.log
for m in ${@:-vbe21_zd ex_zd}; do
  timeout 300 python bench.py --method $m --reads 125000 --steps 3 --warmup 2 --methods "" --no-cpu-baseline --no-e2e > $O/bench_${m}_$TAG.json 2> $O/bench_${m}_$TAG.err || { echo "$m FAILED"; tail -3 $O/bench_${m}_$TAG.err; continue; }
  python -c "
import json; b=json.load(open('$O/bench_${m}_$TAG.json')); print('$m', round(b['compress_gbs']), round(b['decompress_gbs']), {k:round(x['frac'],3) for k,x in b['kernels'].items()})"
done
