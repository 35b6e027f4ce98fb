#!/bin/bash
# GPU parity tests, then the ex-zd exception-rate sweep (20k reads)
O=gpurun_out; TAG=${1:-x}
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest exit $?"; tail -2 $O/pytest_gpu_$TAG.log
bash tools/sweep.sh $TAG 20000
