"""Role timing of the warp-specialised kernels (build with PRESS_GPU_DEFINES=-DPG_ROLE_TIMING).
usage: role_timing.py [method] [nreads]   -> average cycles per tile spent by each role"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from honours_b200 import press
from benchdata import workload

method = sys.argv[1] if len(sys.argv) > 1 else "svb16_zd"
nreads = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
press.init(0)
dev = torch.device("cuda:0")
lengths, x0, shift = workload.read_table(nreads)
off = workload.offsets_of(lengths)
ns = int(off[-1])
d_off = torch.from_numpy(off.astype(np.int64)).to(dev)
d_x0 = torch.from_numpy(x0).to(dev)
d_sh = torch.from_numpy(shift).to(dev)
d_s = torch.empty(ns + 64, dtype=torch.int16, device=dev)
st = torch.cuda.current_stream().cuda_stream
workload.synth_device(d_s.data_ptr(), d_off.data_ptr(), nreads, d_x0.data_ptr(), d_sh.data_ptr(), stream=st)
cap = press.batch_bound(method, nreads, ns)
sb = press.batch_scratch_bytes(method, nreads, ns)
d_out = torch.empty(cap + 64, dtype=torch.uint8, device=dev)
d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
L = press.lib()
L.press_gpu_debug_aux.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
for it in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    press.batch_press(method, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap,
                      d_oo.data_ptr(), d_scr.data_ptr(), sb, st)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
aux = (C.c_uint64 * 29)()
L.press_gpu_debug_aux(d_scr.data_ptr(), st, aux)
a = list(aux)
tiles = ns / 8192 + nreads / 2
sms = torch.cuda.get_device_properties(0).multi_processor_count
comp = int(d_oo[nreads].item())
print(f"{method}: {nreads} reads {ns} samples ~{tiles:.0f} tiles, {ms:.3f} ms (incl. plan), "
      f"{(2 * ns + comp) / ms / 1e6:.0f} GB/s algorithmic; per tile per CTA: {ms * 1e-3 * 1.965e9 / (tiles / sms):.0f} cycles")
rows = [(0, "producer: wait empty", 1), (3, "prefix: wait pub_done", 1), (4, "prefix: info, wait window sum", 1),
        (5, "prefix: P, base, fits, arrive", 1), (21, "window: wait pub_done", 1), (22, "window: sum", 1),
        (23, "window: issue next, hand over", 1), (9, "A warps: wait (empty, full)", 12),
        (10, "A warps: phase A", 12), (15, "B warps: wait scan_done", 12), (16, "B warps: fast blocks", 12),
        (17, "B warps: slow blocks", 12), (24, "prefix: window re-polls (count)", 1)]
for k, nm, n in rows:
    print(f"  {nm:40s} {a[k] / n / tiles:10.1f} per tile per warp")
print(f"  slow blocks: {a[18] / tiles:.3f} per tile, {a[17] / max(a[18], 1):.0f} cycles each; "
      f"fast blocks {a[16] / max(8 * tiles - a[18], 1):.0f} cycles each; phase A {a[10] / (8 * tiles):.0f} cycles per block")
raise SystemExit
for k, nm in enumerate(names):
    if nm != "-":
        print(f"  {nm:34s} {a[k] / nw[k] / tiles:10.1f} per tile per warp")
