TAG=r01g; O=gpurun_out; mkdir -p $O
for m in shuff_vbe21_zd shuff_vbsse21_zd shuff_vbbe21_zd; do
  python bench.py --method $m --reads 125000 --steps 3 --warmup 3 --e2e-gb 1.0 --cpu-reads 1500 --cpu-seconds 4 > $O/bench_${m}_$TAG.json 2> $O/bench_${m}_$TAG.err; echo "$m exit $?"
done
bash tools/ncu_list.sh $TAG 20000 shuff_vbe21_zd > $O/ncu_list_ex_$TAG.log 2>&1
bash tools/ncu_full.sh $TAG 20000 shuff_vbe21_zd huff_decode > $O/ncu_full_shuff_$TAG.log 2>&1
ls $O | grep -c $TAG
