#!/bin/bash
# Huffman emit (tree form): full GPU suite, then the emit variants against the in-tree library, then launch list +
# full captures of the three Huffman kernels with their SASS views (tools/src_lines.py reads them at home)
O=gpurun_out; T=${1:-r02p}; mkdir -p $O
( time timeout 900 python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -8 $O/pytest_gpu_$T.log
M=shuff_vbe21_zd bash tools/r2_var.sh
bash tools/r2_exprof.sh $T shuff_vbe21_zd huff_emit ex_count
