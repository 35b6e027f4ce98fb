"""Summarise `ncu -i X.ncu-rep --page source --csv` (SASS view): per kernel, total
instructions, stall-reason totals, and the hottest instructions by samples."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if len(rows[j]) == len(hdr): body.append(rows[j])
            j += 1
        col = {h: k for k, h in enumerate(hdr)}
        def f(r, h):
            try: return float(r[col[h]])
            except Exception: return 0.0
        tot_inst = sum(f(r, "Instructions Executed") for r in body)
        tot_samp = sum(f(r, "# Samples") for r in body)
        print("=" * 100); print(name); print("warp instructions", int(tot_inst), "samples", int(tot_samp))
        stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        st = {h: sum(f(r, h) for r in body) for h in stalls}
        print("stalls:", ", ".join(f"{h[6:]}={v / max(tot_samp, 1):.1%}" for h, v in sorted(st.items(), key=lambda x: -x[1])[:8]))
        print("shared wavefronts", int(sum(f(r, "L1 Wavefronts Shared") for r in body)), "ideal", int(sum(f(r, "L1 Wavefronts Shared Ideal") for r in body)))
        # how often each SASS line ran (rounded): separates per-read code from the tile loop
        cls = collections.Counter(); cl_n = collections.Counter()
        for r in body:
            e = f(r, "Instructions Executed")
            if e > 0:
                k = float(f"{e:.2g}")
                cls[k] += e; cl_n[k] += 1
        print("by execution count:", ", ".join(f"{int(k)}x: {cl_n[k]} lines = {v / max(tot_inst, 1):.0%}" for k, v in cls.most_common(8)))
        body.sort(key=lambda r: -f(r, "# Samples"))
        for r in body[:top]:
            print(f"{f(r, '# Samples') / max(tot_samp, 1):6.1%} inst={int(f(r, 'Instructions Executed')):>10}  {r[col['Source']][:110]}")
        i = j
    else:
        i += 1
