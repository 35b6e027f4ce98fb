#!/bin/bash
# One gpurun call for the round's evidence: GPU tests, smoke, the reference arm and the default
# bench line (headline + methods + e2e), the remaining methods on one chunk, the ex-zd exception
# sweep, then ncu (launch lists + one full capture per main kernel) on short configurations.
#   gpurun --timeout 2400 -- 'bash tools/gpu_round.sh TAG'
# Everything lands in gpurun_out/ (scratch); tools/round_digest.py TAG copies digests to profiles/.
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/smi_$TAG.csv 2>&1
nproc > $O/nproc_$TAG.txt
python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -2 $O/pytest_gpu_$TAG.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$TAG.log 2>&1; echo "smoke exit $?"; tail -1 $O/smoke_$TAG.log
python bench.py --impl reference > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "ref exit $?"
python bench.py > $O/bench_svb16_zd_$TAG.json 2> $O/bench_svb16_zd_$TAG.err; echo "bench exit $?"; tail -2 $O/bench_svb16_zd_$TAG.err
for m in svb_zd vbz_svb16_zd vbz_svb_zd vbe21_zd vbsse21_zd ex_zd shuff_vbe21_zd shuff_vbsse21_zd vbbe21_zd shuff_vbbe21_zd svb0124_zd uint_zd; do
  python bench.py --method $m --methods "" --reads 125000 --steps 3 --warmup 3 --e2e-gb 1.0 --cpu-reads 1500 --cpu-seconds 4 \
      > $O/bench_${m}_$TAG.json 2> $O/bench_${m}_$TAG.err; echo "$m exit $?"
done
python bench.py --method shuff_vbe21_zd --no-sidecar --methods "" --reads 125000 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline \
    > $O/bench_shuff_vbe21_zd_nosidecar_$TAG.json 2> $O/bench_shuff_vbe21_zd_nosidecar_$TAG.err; echo "no-sidecar exit $?"
bash tools/sweep.sh $TAG 20000 > $O/sweep_$TAG.log 2>&1; tail -6 $O/sweep_$TAG.log
SHORT="python bench.py --reads 8000 --steps 1 --warmup 1 --methods= --no-e2e --no-cpu-baseline"
$SHORT > $O/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/launches_svb16_zd_$TAG.csv $SHORT > $O/ncu_list_$TAG.log 2>&1
echo "ncu list exit $?"
$SHORT > $O/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:svb_ -s 2 -c 2 \
    -f -o $O/prof_svb16_zd_$TAG $SHORT > $O/ncu_full_$TAG.log 2>&1
echo "ncu full exit $?"
# the exception / Huffman methods: launch lists and one full capture per kernel (20 k reads)
bash tools/ncu_list.sh $TAG 20000 vbe21_zd shuff_vbe21_zd svb0124_zd uint_zd > $O/ncu_list_ex_$TAG.log 2>&1
bash tools/ncu_full.sh $TAG 20000 vbe21_zd ex_count ex_emit_bytes ex_decode2 > $O/ncu_full_vbe21_$TAG.log 2>&1
bash tools/ncu_full.sh $TAG 20000 shuff_vbe21_zd ex_count huff_emit huff_fused > $O/ncu_full_shuff_$TAG.log 2>&1
bash tools/ncu_full.sh $TAG 20000 svb0124_zd mc_size mc_encode mc_decode > $O/ncu_full_svb0124_$TAG.log 2>&1
# digest on the box (ncu is here), then drop the captures: gpurun brings back at most 64 MiB
python tools/round_digest.py $TAG $O/digest_$TAG > $O/digest_$TAG.log 2>&1; tail -2 $O/digest_$TAG.log
rm -f $O/*.ncu-rep
ls $O | grep -c $TAG
