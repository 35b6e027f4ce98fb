#!/bin/bash
# after a change to shared helpers: GPU suite, then short bench lines of a few methods (125 k reads) and the sweep points
O=gpurun_out; T=${1:-r02d}; shift; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gpu_$T.log
for m in "$@"; do
python bench.py --method $m --methods "" --reads 125000 --steps 3 --warmup 2 --no-e2e --no-cpu-baseline > $O/bench_${m}_$T.json 2> $O/bench_${m}_$T.err
python -c "
import json; b=json.load(open('$O/bench_${m}_$T.json')); print('$m', round(b['value']), {k:(round(v['frac'],3),round(v['ms_per_launch'],2)) for k,v in b['kernels'].items()})"
done
for p in 1e-3 1e-2; do
python bench.py --method ex_zd --reads 20000 --p-ex $p --steps 2 --warmup 1 --methods= --no-e2e --no-cpu-baseline > $O/sw_${p}_$T.json 2>/dev/null
python -c "
import json; b=json.load(open('$O/sw_${p}_$T.json')); print('ex_zd p_ex=$p', {k:round(v['frac'],3) for k,v in b['kernels'].items()})"
done
