"""The measured-numbers table of BASELINE.md section 5 from the bench lines of one evidence round:
    python tools/baseline_table.py r02b        (reads profiles/r02b_<method>_bench.json)"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
ROWS = [("svb16_zd", "svb16-zd"), ("svb_zd", "svb-zd"), ("vbz_svb16_zd", "VBZ frame, svb16"), ("vbz_svb_zd", "VBZ frame, svb"),
        ("vbe21_zd", "vbe21-zd"), ("vbsse21_zd", "vbsse21-zd"), ("vbbe21_zd", "vbbe21-zd"), ("ex_zd", "ex-zd"),
        ("shuff_vbe21_zd", "shuff-vbe21-zd"), ("shuff_vbsse21_zd", "shuff-vbsse21-zd"), ("shuff_vbbe21_zd", "shuff-vbbe21-zd"),
        ("shuff_vbe21_zd_nosidecar", "shuff-vbe21-zd, no side-car"), ("svb0124_zd", "svb0124-zd"), ("uint_zd", "uint-zd")]
print("| method | B/sample | compress | decompress | round trip (`value`) | encode / decode % of HBM | e2e via host buffers (duplex) | reference code, host threads of the box |")
print("|---|---|---|---|---|---|---|---|")
for m, name in ROWS:
    f = os.path.join(ROOT, "profiles", f"{tag}_{m}_bench.json")
    if not os.path.exists(f):
        continue
    d = json.load(open(f))
    k = d["kernels"]
    e2e = d.get("e2e") or {}
    cb = d.get("cpu_baseline") or {}
    val = "**%.0f**" % d["value"] if m == "svb16_zd" else "%.0f" % d["value"]
    print("| %s | %.4f | %.0f | %.0f | %s | %.1f / %.1f | %s | %s |" % (
        name, d["bytes_per_sample"], d["compress_gbs"], d["decompress_gbs"], val, 100 * k["encode"]["frac"], 100 * k["decode"]["frac"],
        "%.1f" % e2e["value"] if e2e.get("value") else "-",
        "%.1f (%d threads)" % (cb["value"], cb["cores"]) if cb.get("value") else "-"))
