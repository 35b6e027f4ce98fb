#!/bin/bash
# variants only (decode): tools/r2_run11.sh METHOD v1 v2 ...
M=$1; shift
mkdir -p gpurun_out
M=$M bash tools/r2_var.sh "$@"
