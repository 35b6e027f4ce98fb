#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, then ncu (launch list + one
# full capture of the codec kernels) on a short configuration of the same bench.
#   gpurun --timeout 1700 -- 'bash tools/gpu_round.sh [method] [tag]'
# Everything lands in gpurun_out/ (scratch); copy what should be judged into profiles/.
METHOD=${1:-svb16_zd}
TAG=${2:-r1}
KREGEX=${3:-svb_zd}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/smi_$TAG.csv 2>&1
python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1
echo "pytest exit $?" | tee -a $O/pytest_gpu_$TAG.log
tail -3 $O/pytest_gpu_$TAG.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$TAG.log 2>&1
echo "smoke exit $?" | tee -a $O/smoke_$TAG.log
python bench.py --method $METHOD --steps 3 --warmup 3 > $O/bench_${METHOD}_$TAG.json 2> $O/bench_${METHOD}_$TAG.err
echo "bench exit $?"
cat $O/bench_${METHOD}_$TAG.json
SHORT="python bench.py --method $METHOD --reads 5000 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
$SHORT > $O/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $O/launches_${METHOD}_$TAG.csv $SHORT > $O/ncu_list_$TAG.log 2>&1
echo "ncu list exit $?"
$SHORT > $O/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 2 -c 2 \
    -f -o $O/prof_${METHOD}_$TAG $SHORT > $O/ncu_full_$TAG.log 2>&1
echo "ncu full exit $?"
ls -la $O
