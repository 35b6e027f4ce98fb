#!/bin/bash
O=gpurun_out; T=${1:-r02c}
mkdir -p $O
( time timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sidecar or huffman or bench_scale" ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -15 $O/pytest_gpu_$T.log
timeout 300 python bench.py --method shuff_vbe21_zd --reads 125000 --steps 3 --warmup 2 --methods "" --no-cpu-baseline --no-e2e > $O/bench_shuff_$T.json 2> $O/bench_shuff_$T.err; echo "bench exit $?"; tail -3 $O/bench_shuff_$T.err
python -c "
import json; b=json.load(open('$O/bench_shuff_$T.json')); print(b['value'], b['compress_gbs'], b['decompress_gbs'], {k:(v['frac'],v['ms_per_launch']) for k,v in b['kernels'].items()})"
