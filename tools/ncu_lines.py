"""Per CUDA source line: instructions executed and stall samples, from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
fn = None; path = None; col = None; agg = {}
order = []
for r in rows:
    if not r: continue
    if r[0] == "File Path": path = r[1].split("/")[-1]; continue
    if r[0] == "Function Name":
        fn = r[1]
        if fn not in agg: agg[fn] = {}; order.append(fn)
        continue
    if r[0] == "Line No": col = {h: k for k, h in enumerate(r)}; iI = r.index("Instructions Executed"); iS = r.index("# Samples"); continue
    if col and r[0] not in ("", "-") and r[0].isdigit():
        key = (path, int(r[0]), r[1].strip()[:90])
        try:
            a = agg[fn].setdefault(key, [0.0, 0.0]); a[0] += float(r[iI]); a[1] += float(r[iS])
        except ValueError: pass
seen = set()
for f in order:
    if f in seen: continue
    seen.add(f)
    d = agg[f]; ti = sum(v[0] for v in d.values()); ts = sum(v[1] for v in d.values())
    print("=" * 90); print(f, "inst", int(ti), "samples", int(ts))
    for k, v in sorted(d.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{v[0] / ti:6.1%} inst {v[1] / max(ts, 1):6.1%} smp  {k[0]}:{k[1]:<4} {k[2]}")
