#!/bin/bash
# BASELINE.json config 3: ex_zd exception-rate sweep (device-resident, one chunk): tools/sweep.sh TAG READS
TAG=${1:-r01b}; READS=${2:-20000}
mkdir -p gpurun_out
for p in 0 4.28e-5 1e-3 1e-2 1e-1 0.5; do
  timeout 900 python bench.py --method ex_zd --reads $READS --p-ex $p --steps 2 --warmup 1 --methods= --no-e2e --no-cpu-baseline \
      > gpurun_out/sweep_ex_zd_${p}_$TAG.json 2> gpurun_out/sweep_ex_zd_${p}_$TAG.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/sweep_ex_zd_${p}_$TAG.json"))
    k = d["kernels"]
    print("ex_zd p_ex=$p", "B/sample %.4f" % d["bytes_per_sample"], "compress %.0f decompress %.0f GB/s raw" % (d["compress_gbs"], d["decompress_gbs"]),
          {n: "%.1f%% of HBM peak" % (100 * v["achieved_gbs"] / d["roofline"]["peak"]) for n, v in k.items()})
except Exception as e:
    print("ex_zd p_ex=$p: no result", e); print(open("gpurun_out/sweep_ex_zd_${p}_$TAG.err").read()[-400:])
PY
done
