#!/bin/bash
# Refresh of one evidence round for the methods that use the bit packer (after a change to it):
# bench lines of the Huffman and svb0124 / uint methods, captures of the Huffman encode kernels.  tools/gpu_round_huff.sh TAG
TAG=${1:-r02c}; O=gpurun_out; mkdir -p $O
for m in shuff_vbe21_zd shuff_vbsse21_zd shuff_vbbe21_zd svb0124_zd uint_zd; do
  python bench.py --method $m --methods "" --reads 125000 --steps 3 --warmup 3 --e2e-gb 1.0 --cpu-reads 1500 --cpu-seconds 4 \
      > $O/bench_${m}_$TAG.json 2> $O/bench_${m}_$TAG.err; echo "$m exit $?"
done
bash tools/r2_exprof.sh $TAG shuff_vbe21_zd ex_count huff_emit | grep -E "ncu exit|time_duration"
