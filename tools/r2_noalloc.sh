#!/bin/bash
# the no-allocation entry point's test: alone in a fresh process, then after every other GPU test
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone exit $?"; tail -5 gpurun_out/pytest_noalloc_alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_noalloc_all.log 2>&1; echo "all exit $?"; tail -4 gpurun_out/pytest_gpu_noalloc_all.log
