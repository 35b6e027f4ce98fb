#!/bin/bash
# Huffman fused decode: long codes inline / 14-bit first level; full GPU suite, variants, capture with SASS view
O=gpurun_out; T=${1:-r02t}; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -8 $O/pytest_gpu_$T.log
M=shuff_vbe21_zd bash tools/r2_var.sh "$@"
bash tools/r2_exprof.sh $T shuff_vbe21_zd huff_fused
