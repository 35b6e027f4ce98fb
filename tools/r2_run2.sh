#!/bin/bash
O=gpurun_out; T=${1:-r02b}
mkdir -p $O
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu_$T.log
python bench.py --steps 3 --warmup 2 --methods "" --no-cpu-baseline > $O/bench_$T.json 2> $O/bench_$T.err; echo "bench exit $?"; tail -3 $O/bench_$T.err
python -c "
import json; b=json.load(open('$O/bench_$T.json')); print('e2e', b['e2e']['value'], b['e2e']['serial_value'], b['e2e']['link_ceiling_gbs'])"
