"""debug: device batch API vs oracle on a random batch; first-difference diagnostics.
usage: debug_batch.py [method] [nreads] [lo] [hi] [--nodec]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from honours_b200 import press
from oracle import cpu_impl


def main():
    method = sys.argv[1] if len(sys.argv) > 1 else "svb16_zd"
    nreads = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
    lo = int(sys.argv[3]) if len(sys.argv) > 3 else 2000
    hi = int(sys.argv[4]) if len(sys.argv) > 4 else 60000
    cpu_impl.build(ref=False)
    o = cpu_impl.load("oracle")
    tab = open(os.path.join(os.path.dirname(press.__file__), "data", "NA12878_zd.huffman"), "rb").read()
    o.set_huffman_table(tab)
    press.init(0)
    if method == "shuff_vbe21_zd":
        press.set_huffman_table(tab)
    rng = np.random.default_rng(5)
    lens = rng.integers(lo, hi, nreads)
    reads = []
    for i, n in enumerate(lens):
        d = rng.integers(-12, 13, int(n))
        if i % 2:
            idx = rng.random(int(n)) < 0.01
            d[idx] = rng.integers(-900, 900, int(idx.sum()))
        reads.append((500 + np.cumsum(d)).astype(np.int16))
    samples = np.concatenate(reads)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    dev = torch.device("cuda:0")
    d_s = torch.from_numpy(samples).to(dev)
    d_off = torch.from_numpy(offsets.astype(np.int64)).to(dev)
    ns = samples.size
    cap = press.batch_bound(method, nreads, ns)
    d_out = torch.zeros(cap, dtype=torch.uint8, device=dev)
    d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
    sb = press.batch_scratch_bytes(method, nreads, ns)
    d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    press.batch_press(method, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap,
                      d_oo.data_ptr(), d_scr.data_ptr(), sb, st)
    print("status", press.batch_status(d_scr.data_ptr(), st), press.last_error())
    oo = d_oo.cpu().numpy()
    total = int(oo[-1])
    expect = [o.press(method, s)[1] for s in reads]
    exp = np.frombuffer(b"".join(expect), np.uint8)
    got = d_out[:total].cpu().numpy()
    eo = np.concatenate([[0], np.cumsum([len(e) for e in expect])])
    print("total", total, "expect", exp.size, "offsets equal", np.array_equal(oo, eo))
    mm = min(got.size, exp.size)
    d = np.nonzero(got[:mm] != exp[:mm])[0]
    print("ndiff", d.size)
    if d.size:
        for x in d[:10]:
            r = int(np.searchsorted(eo, x, side="right") - 1)
            print(f"  byte {x} read {r} (+{x - eo[r]} of {eo[r + 1] - eo[r]}, n={lens[r]}) got {got[x]} want {exp[x]}")
    if "--nodec" in sys.argv:
        return
    d_back = torch.zeros(ns, dtype=torch.int16, device=dev)
    d_in = torch.from_numpy(exp.copy()).to(dev)
    d_eo = torch.from_numpy(eo.astype(np.int64)).to(dev)
    press.batch_depress(method, d_in.data_ptr(), d_eo.data_ptr(), exp.size, d_off.data_ptr(), nreads, ns,
                        d_back.data_ptr(), d_scr.data_ptr(), sb, st)
    print("dec status", press.batch_status(d_scr.data_ptr(), st), press.last_error())
    bad = (d_back != d_s).nonzero().flatten().cpu().numpy()
    print("decode mismatches", bad.size, bad[:10])


main()
