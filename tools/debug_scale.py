"""debug: find where the round trip breaks at large chunk sizes"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from honours_b200 import press
from benchdata import workload

def run(nreads, method="svb16_zd"):
    dev = torch.device("cuda:0")
    press.init(0)
    lengths, x0, shift = workload.read_table(nreads)
    off = workload.offsets_of(lengths)
    ns = int(off[-1])
    d_off = torch.from_numpy(off.astype(np.int64)).to(dev)
    d_x0 = torch.from_numpy(x0).to(dev); d_sh = torch.from_numpy(shift).to(dev)
    d_s = torch.empty(ns + 64, dtype=torch.int16, device=dev)
    d_back = torch.zeros(ns + 64, dtype=torch.int16, device=dev)
    cap = press.batch_bound(method, nreads, ns)
    sb = press.batch_scratch_bytes(method, nreads, ns)
    d_out = torch.empty(cap + 64, dtype=torch.uint8, device=dev)
    d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
    d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    workload.synth_device(d_s.data_ptr(), d_off.data_ptr(), nreads, d_x0.data_ptr(), d_sh.data_ptr(), stream=st)
    press.batch_press(method, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap, d_oo.data_ptr(), d_scr.data_ptr(), sb, st)
    print("press status", press.batch_status(d_scr.data_ptr(), st))
    oo = d_oo.cpu().numpy()
    print("ns", ns, "total", oo[-1], "B/sample", oo[-1] / ns, "monotone", bool((np.diff(oo) > 0).all()))
    # expected size per read from keys: nkeys + n + popcount(keys)
    press.batch_depress(method, d_out.data_ptr(), d_oo.data_ptr(), cap, d_off.data_ptr(), nreads, ns, d_back.data_ptr(), d_scr.data_ptr(), sb, st)
    print("depress status", press.batch_status(d_scr.data_ptr(), st))
    step = 1 << 28
    nbad = 0; first = None
    for i in range(0, ns, step):
        ne = (d_s[i:min(ns, i + step)] != d_back[i:min(ns, i + step)])
        c = int(ne.sum().item())
        if c and first is None:
            first = i + int(ne.nonzero()[0].item())
        nbad += c
    print("mismatching samples", nbad, "first", first)
    if first is not None:
        r = int(np.searchsorted(off, first, side="right") - 1)
        print("read", r, "off", off[r], "n", lengths[r], "pos in read", first - int(off[r]), "tile", (first - int(off[r])) // 4096)
        print("orig", d_s[first - 4:first + 12].cpu().numpy()); print("back", d_back[first - 4:first + 12].cpu().numpy())
        # count bad per read for first few bad reads
        ne_idx = []
        # check the oracle on that read
        from oracle import cpu_impl
        cpu_impl.build(ref=False); o = cpu_impl.load("oracle")
        s = d_s[int(off[r]):int(off[r + 1])].cpu().numpy()
        rc, stt = o.press(method, s)
        got = d_out[int(oo[r]):int(oo[r + 1])].cpu().numpy().tobytes()
        print("stream equal oracle:", got == stt, len(got), len(stt))
        if got != stt:
            a = np.frombuffer(got, np.uint8); b = np.frombuffer(stt, np.uint8)
            m = min(a.size, b.size); d = np.nonzero(a[:m] != b[:m])[0]
            print("first diff byte", d[:5], "of", m, "nkeys", (s.size + 7) // 8)

for nr in [int(x) for x in sys.argv[1:]]:
    print("=== nreads", nr); run(nr); torch.cuda.empty_cache()
