"""debug: per-read press/depress vs oracle with first-difference diagnostics"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from honours_b200 import press
from oracle import cpu_impl

def main():
    methods = sys.argv[1].split(",") if len(sys.argv) > 1 else press.implemented_methods()
    cpu_impl.build(ref=False)
    o = cpu_impl.load("oracle")
    tab = open(os.path.join(os.path.dirname(press.__file__), "data", "NA12878_zd.huffman"), "rb").read()
    o.set_huffman_table(tab)
    press.init(0)
    if "shuff_vbe21_zd" in methods:
        press.set_huffman_table(tab)
    rng = np.random.default_rng(1)
    cases = []
    for n in [1, 5, 8, 9, 100, 1023, 1024, 1025, 2048, 5000, 70000]:
        cases.append(("walk", n, (475 + np.cumsum(rng.integers(-25, 26, n))).astype(np.int16)))
        d = rng.integers(-12, 13, n); idx = rng.random(n) < 0.01; d[idx] = rng.integers(-900, 900, int(idx.sum()))
        cases.append(("spiky", n, (500 + np.cumsum(d)).astype(np.int16)))
        cases.append(("noise", n, rng.integers(-32768, 32768, n).astype(np.int16)))
        cases.append(("q", n, ((100 + np.cumsum(rng.integers(-6, 7, n))) << 2).astype(np.int16)))
    nfail = 0
    for method in methods:
        for kind, n, s in cases:
            rc, st = press.press(method, s)
            ro, so = o.press(method, s)
            ok = rc == ro and (ro != 0 or st == so)
            msg = ""
            if not ok:
                a = np.frombuffer(st, np.uint8); b = np.frombuffer(so, np.uint8)
                mm = min(a.size, b.size); d = np.nonzero(a[:mm] != b[:mm])[0]
                msg = f"PRESS rc {rc}/{ro} len {a.size}/{b.size} firstdiff {d[:4]} got {a[d[:4]] if d.size else ''} want {b[d[:4]] if d.size else ''}"
            if ro == 0:
                rc2, back = press.depress(method, so, n)
                if rc2 != 0 or not np.array_equal(back, s):
                    d = np.nonzero(back[:n] != s[:back.size])[0] if back.size else np.array([])
                    msg += f" DEPRESS rc {rc2} n {back.size}/{n} firstdiff {d[:4]} got {back[d[:4]] if d.size else ''} want {s[d[:4]] if d.size else ''} ndiff {d.size}"
                    ok = False
            if not ok:
                nfail += 1
                print(method, kind, n, msg)
    print("failures:", nfail)

main()
