"""Digest a gpurun ncu capture into the small text files kept under profiles/.

    python tools/profile_digest.py <tag> <launches.csv> <prof.ncu-rep> [bench.json]

writes profiles/<tag>_launches.txt   (per-kernel launch count / total / share of GPU time)
       profiles/<tag>_kernels.txt    (key raw metrics + stall reasons + hottest SASS per kernel)
and copies the bench line to profiles/<tag>_bench.json when given.
"""
import collections
import csv
import io
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "lts__t_sector_hit_rate.pct", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__shared_mem_per_block_static",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__cycles_elapsed.max",
]


def launches(path, out):
    rows = [r for r in csv.reader(open(path)) if len(r) > 14 and r[0].isdigit()]
    agg = collections.OrderedDict()
    for r in rows:
        name = r[4].split("(")[0].replace("void ", "")
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += float(r[14])
    tot = sum(v[1] for v in agg.values()) or 1.0
    with open(out, "w") as f:
        f.write(f"# {os.path.basename(path)}: {len(rows)} launches, {tot / 1e6:.3f} ms of GPU time "
                "(ncu gpu__time_duration.sum, --clock-control none; cold-cache, serialised: compare shares)\n")
        f.write(f"{'kernel':70s} {'n':>5s} {'total ms':>10s} {'avg us':>10s} {'share':>7s}\n")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{k[:70]:70s} {n:5d} {t / 1e6:10.3f} {t / n / 1e3:10.1f} {t / tot:7.1%}\n")


def kernels(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    tmp = "/tmp/_src.csv"
    open(tmp, "w").write(src)
    summ = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), tmp, "14"],
                          capture_output=True, text=True).stdout
    with open(out, "w") as f:
        f.write(f"# {os.path.basename(rep)}: ncu --set full --clock-control none (per launch)\n")
        for r in rows[2:]:
            f.write("\n== " + r[hdr.index("Kernel Name")] + "\n")
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    f.write(f"{k:62s} {r[i]:>16s} {units[i]}\n")
            try:
                rd = float(r[hdr.index("dram__bytes_read.sum")])
                wr = float(r[hdr.index("dram__bytes_write.sum")])
                t = float(r[hdr.index("gpu__time_duration.sum")])
                uu = [units[hdr.index(x)] for x in ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum")]
                f.write(f"{'traffic (read+write)':62s} {rd + wr:16.6f} {uu[0]}   per {t} {uu[2]}\n")
            except Exception:
                pass
        f.write("\n# stall reasons / hottest SASS (ncu --page source), one block per captured launch\n")
        seen = set()
        for blk in summ.split("=" * 100):
            key = blk.strip()[:4000]
            if key and key not in seen:
                seen.add(key)
                f.write("=" * 100 + blk)


if __name__ == "__main__":
    tag, lcsv, rep = sys.argv[1:4]
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    launches(lcsv, os.path.join(ROOT, "profiles", f"{tag}_launches.txt"))
    kernels(rep, os.path.join(ROOT, "profiles", f"{tag}_kernels.txt"))
    if len(sys.argv) > 4:
        shutil.copy(sys.argv[4], os.path.join(ROOT, "profiles", f"{tag}_bench.json"))
    print("wrote profiles/" + tag + "_*")
