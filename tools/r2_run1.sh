#!/bin/bash
# round 2, GPU call 1: infrastructure check -- GPU tests (now against oracle/_ref), smoke, both bench arms
O=gpurun_out; T=${1:-r02a}
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/smi_$T.csv 2>&1
nproc > $O/nproc_$T.txt
( time python -m pytest tests -m gpu -x -q ) > $O/pytest_gpu_$T.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu_$T.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_$T.log 2>&1; echo "smoke exit $?"; tail -1 $O/smoke_$T.log
( time python bench.py --impl reference --steps 3 --warmup 1 ) > $O/bench_ref_$T.json 2> $O/bench_ref_$T.err; echo "ref exit $?"
( time python bench.py --steps 5 --warmup 3 ) > $O/bench_$T.json 2> $O/bench_$T.err; echo "bench exit $?"; tail -5 $O/bench_$T.err
head -c 600 $O/bench_$T.json
