"""Instructions executed per CUDA source line: joins the SASS view of a capture
(ncu -i X.ncu-rep --page source --csv, saved by tools/r2_exprof.sh) with the line table of the same
build (nvdisasm -g on the cubin of libpress_gpu.so).
    python tools/src_lines.py gpurun_out/sass_M_K_TAG.csv <kernel substring> [top]"""
import collections, csv, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 45
lib = os.path.join(ROOT, "honours_b200", "libpress_gpu.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, check=True, capture_output=True)
line_of = {}
for cub in os.listdir(tmp):
    dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
    infn = False; cur = None
    for ln in dis.split("\n"):
        m = re.match(r"\s*\.section\s+\.text\.(\S+),", ln)
        if m:
            infn = kern in m.group(1); cur = None; continue
        if not infn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+\S", ln)
        if m and cur:
            line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(path)))
hi = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
hdr = rows[hi]; c = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
acol = c.get("Address")
agg = collections.Counter(); smp = collections.Counter()
base = None
for k, r in enumerate(body):
    try:
        e = float(r[c["Instructions Executed"]]); s = float(r[c["# Samples"]])
    except ValueError:
        continue
    if acol is not None:
        a = int(r[acol], 16) if r[acol].startswith("0x") else int(r[acol])
        if base is None: base = a
        off = a - base
    else:
        off = 16 * k
    key = line_of.get(off, ("?", 0))
    agg[key] += e; smp[key] += s
tot = sum(agg.values()) or 1; ts = sum(smp.values()) or 1
src = {}
def text(f, n):
    if f not in src:
        p = os.path.join(ROOT, "honours_b200", "csrc", f)
        src[f] = open(p).read().split("\n") if os.path.exists(p) else []
    return src[f][n - 1].strip()[:100] if 0 < n <= len(src[f]) else ""
print("warp instructions", int(tot), "mapped lines", len(agg))
for (f, n), e in agg.most_common(top):
    print(f"{e / tot:6.1%} inst {smp[(f, n)] / ts:6.1%} samp  {f}:{n:<5} {text(f, n)}")
