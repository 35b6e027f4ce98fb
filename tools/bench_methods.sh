#!/bin/bash
# short device-resident bench of several methods (no e2e / CPU legs): tools/bench_methods.sh TAG READS m1 m2 ...
TAG=$1; READS=$2; shift 2
mkdir -p gpurun_out
for m in "$@"; do
  timeout 600 python bench.py --method $m --reads $READS --steps 2 --warmup 1 --no-e2e --no-cpu-baseline \
      > gpurun_out/bm_${m}_$TAG.json 2> gpurun_out/bm_${m}_$TAG.err
  echo "$m exit $?"; tail -2 gpurun_out/bm_${m}_$TAG.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bm_${m}_$TAG.json"))
    k = d["kernels"]
    print("$m", "value %.0f" % d["value"], "comp %.0f decomp %.0f GB/s raw" % (d["compress_gbs"], d["decompress_gbs"]),
          "B/sample %.4f" % d["bytes_per_sample"],
          {n: "%.0f GB/s moved (%.1f%%)" % (v["achieved_gbs"], 100 * v["achieved_gbs"] / d["roofline"]["peak"]) for n, v in k.items()})
except Exception as e:
    print("$m: no result", e)
PY
done
