"""Turn one tools/gpu_round.sh run (gpurun_out/*_TAG*) into the files kept under profiles/:
bench lines, launch lists, per-kernel ncu digests, the exception sweep and traffic.json.
    python tools/round_digest.py TAG"""
import json
import os
import re
import shutil
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import profile_digest as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
tag = sys.argv[1]
# (on the GPU box the digests go to a directory under gpurun_out/, the only one that travels back;
#  the .ncu-rep files themselves are too large to bring home)
P = os.path.abspath(sys.argv[2]) if len(sys.argv) > 2 else os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)
METHODS = ["svb16_zd", "svb_zd", "vbz_svb16_zd", "vbz_svb_zd", "vbe21_zd", "vbsse21_zd", "ex_zd", "shuff_vbe21_zd",
           "shuff_vbsse21_zd", "vbbe21_zd", "shuff_vbbe21_zd", "svb0124_zd", "uint_zd", "shuff_vbe21_zd_nosidecar"]


def last_json(path):
    return json.loads([ln for ln in open(path).read().splitlines() if ln.startswith("{")][-1])


for m in METHODS:
    src = os.path.join(G, f"bench_{m}_{tag}.json")
    if os.path.exists(src):
        json.dump(last_json(src), open(os.path.join(P, f"{tag}_{m}_bench.json"), "w"))
for a, b in [(f"bench_ref_{tag}.json", f"{tag}_reference_arm_bench.json")]:
    if os.path.exists(os.path.join(G, a)):
        json.dump(last_json(os.path.join(G, a)), open(os.path.join(P, b), "w"))
for a, b in [(f"pytest_gpu_{tag}.log", f"{tag}_pytest_gpu.log"), (f"smoke_{tag}.log", f"{tag}_smoke.log")]:
    if os.path.exists(os.path.join(G, a)):
        shutil.copy(os.path.join(G, a), os.path.join(P, b))

pd.launches(os.path.join(G, f"launches_svb16_zd_{tag}.csv"), os.path.join(P, f"{tag}_svb16_zd_launches.txt"))
pd.kernels(os.path.join(G, f"prof_svb16_zd_{tag}.ncu-rep"), os.path.join(P, f"{tag}_svb16_zd_kernels.txt"))
caps = {"vbe21_zd": ["ex_count", "ex_emit_bytes", "ex_decode2"],
        "shuff_vbe21_zd": ["ex_count", "huff_emit", "huff_fused"],
        "svb0124_zd": ["mc_size", "mc_encode", "mc_decode"], "uint_zd": []}
for m, ks in caps.items():
    lc = os.path.join(G, f"launches_{m}_{tag}.csv")
    if os.path.exists(lc):
        pd.launches(lc, os.path.join(P, f"{tag}_{m}_launches.txt"))
    for k in ks:
        rep = os.path.join(G, f"full_{m}_{k}_{tag}.ncu-rep")
        if os.path.exists(rep):
            pd.kernels(rep, os.path.join(P, f"{tag}_{m}_{k}_kernels.txt"))

rows = []
for pe in ["0", "4.28e-5", "1e-3", "1e-2", "1e-1", "0.5"]:
    f = os.path.join(G, f"sweep_ex_zd_{pe}_{tag}.json")
    if os.path.exists(f):
        d = last_json(f)
        rows.append({"p_ex": pe, "bytes_per_sample": d["bytes_per_sample"], "compress_gbs_raw": d["compress_gbs"],
                     "decompress_gbs_raw": d["decompress_gbs"],
                     "encode_frac_of_hbm_peak": d["kernels"]["encode"]["achieved_gbs"] / d["roofline"]["peak"],
                     "decode_frac_of_hbm_peak": d["kernels"]["decode"]["achieved_gbs"] / d["roofline"]["peak"]})
if rows:
    json.dump({"workload": "ex_zd, 20000 synthetic reads (2.3 G samples) device resident, bench.py --method ex_zd "
                           "--reads 20000 --p-ex P --steps 2 --warmup 1", "rows": rows},
              open(os.path.join(P, f"{tag}_ex_zd_exception_sweep.json"), "w"), indent=1)

# ---- traffic.json: DRAM bytes of each method's encode / decode kernel sequence, per launch
MUL = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def dram(path, kernel=None):
    t = open(path).read()
    if kernel:
        t = t[t.index(kernel):]
    tot = 0.0
    for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        mo = re.search(key + r"\s+([\d.]+) (\w+)", t)
        tot += float(mo.group(1)) * MUL[mo.group(2)]
    return tot


traffic = {}
NS20K = 2305293902          # samples of `--reads 20000`
NS8K = None
try:
    b = last_json(os.path.join(G, f"plain_{tag}.log"))
    NS8K = b["config"]["samples_per_gpu"]
    alg8 = b["kernels"]["encode"]["alg_bytes_per_launch"]
    kf = os.path.join(P, f"{tag}_svb16_zd_kernels.txt")
    for nm, kname in (("encode", "svb_enc_kernel"), ("decode", "svb_dec2_kernel")):
        traffic[f"svb16_zd_{nm}"] = {"dram_bytes": dram(kf, kname), "alg_bytes": alg8,
                                     "source": f"profiles/{tag}_svb16_zd_kernels.txt (ncu --set full, bench.py --reads 8000, {kname}<1,0>)"}
except Exception as e:
    print("svb16 traffic:", e)
for m, enc, dec, bps in (("vbe21_zd", ["ex_count", "ex_emit_bytes"], ["ex_decode2"], 1.0002),
                         ("shuff_vbe21_zd", ["ex_count", "huff_emit"], ["huff_fused"], 0.6788)):
    try:
        alg = (2.0 + bps) * NS20K
        for nm, ks in (("encode", enc), ("decode", dec)):
            tot = sum(dram(os.path.join(P, f"{tag}_{m}_{k}_kernels.txt")) for k in ks)
            traffic[f"{m}_{nm}"] = {"dram_bytes": tot, "alg_bytes": alg,
                                    "source": f"profiles/{tag}_{m}_*_kernels.txt (ncu --set full, bench.py --method {m} --reads 20000; "
                                              + " + ".join(ks) + ")"}
    except Exception as e:
        print(m, "traffic:", e)
# same kernels as the captured method; only the header coding differs
for base, twins in (("vbe21_zd", ("vbsse21_zd", "vbbe21_zd", "ex_zd")), ("shuff_vbe21_zd", ("shuff_vbsse21_zd", "shuff_vbbe21_zd"))):
    for twin in twins:
        for nm in ("encode", "decode"):
            if f"{base}_{nm}" in traffic:
                traffic[f"{twin}_{nm}"] = dict(traffic[f"{base}_{nm}"],
                                               source=traffic[f"{base}_{nm}"]["source"] + f" [{base} capture: same kernels]")
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
print("profiles/ updated for", tag, "; traffic keys:", sorted(traffic))
