#!/bin/bash
# diagnostic: same bench with read lengths clipped (no long-read tail): tools/bench_cap.sh TAG READS CAP m1 m2 ...
TAG=$1; READS=$2; CAP=$3; shift 3
for m in "$@"; do
  timeout 600 python bench.py --method $m --reads $READS --len-cap $CAP --steps 2 --warmup 1 --no-e2e --no-cpu-baseline \
      > gpurun_out/cap_${m}_$TAG.json 2> gpurun_out/cap_${m}_$TAG.err
  python - <<PY
import json
d = json.load(open("gpurun_out/cap_${m}_$TAG.json"))
print("$m cap $CAP", "comp %.0f decomp %.0f GB/s raw" % (d["compress_gbs"], d["decompress_gbs"]),
      {n: "%.0f GB/s moved (%.1f%%)" % (v["achieved_gbs"], 100 * v["achieved_gbs"] / d["roofline"]["peak"]) for n, v in d["kernels"].items()})
PY
done
