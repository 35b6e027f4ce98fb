#!/bin/bash
# svb0124-zd / uint-zd through staged tiles and the Huffman emitter's bit packer: GPU suite, then the two bench lines
O=gpurun_out; T=${1:-r02u}; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -25 $O/pytest_gpu_$T.log
for m in svb0124_zd uint_zd; do
timeout 300 python bench.py --method $m --reads 60000 --steps 2 --warmup 1 --methods "" --no-cpu-baseline --no-e2e > $O/bench_${m}_$T.json 2> $O/bench_${m}_$T.err; echo "bench exit $?"; tail -3 $O/bench_${m}_$T.err
python -c "
import json; b=json.load(open('$O/bench_${m}_$T.json')); print('$m', b['value'], b['compress_gbs'], b['decompress_gbs'], b['bytes_per_sample'], {k:(v['frac'],v['ms_per_launch']) for k,v in b['kernels'].items()})"
done
