"""Huffman decode statistics: steps and re-synchronisation rounds per step (hdr aux words)."""
import ctypes as C, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from honours_b200 import press
from benchdata import workload
press.init(0); press.set_huffman_table()
nreads = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
dev = torch.device("cuda:0")
lengths, x0, shift = workload.read_table(nreads)
off = workload.offsets_of(lengths); ns = int(off[-1])
d_off = torch.from_numpy(off.astype(np.int64)).to(dev)
d_x0 = torch.from_numpy(x0).to(dev); d_sh = torch.from_numpy(shift).to(dev)
d_s = torch.empty(ns + 64, dtype=torch.int16, device=dev)
st = torch.cuda.current_stream().cuda_stream
workload.synth_device(d_s.data_ptr(), d_off.data_ptr(), nreads, d_x0.data_ptr(), d_sh.data_ptr(), stream=st)
m = "shuff_vbe21_zd"
cap = int(1.2 * ns) + 4096; sb = press.batch_scratch_bytes(m, nreads, ns)
d_out = torch.empty(cap, dtype=torch.uint8, device=dev); d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
d_scr = torch.empty(sb, dtype=torch.uint8, device=dev); d_back = torch.empty(ns + 64, dtype=torch.int16, device=dev)
press.batch_press(m, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap, d_oo.data_ptr(), d_scr.data_ptr(), sb, st)
assert press.batch_status(d_scr.data_ptr(), st) == 0
for it in range(2):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    press.batch_depress(m, d_out.data_ptr(), d_oo.data_ptr(), cap, d_off.data_ptr(), nreads, ns, d_back.data_ptr(), d_scr.data_ptr(), sb, st)
    e1.record(); torch.cuda.synchronize()
aux = (C.c_uint64 * 29)()
press.lib().press_gpu_debug_aux(C.c_void_p(d_scr.data_ptr()), C.c_void_p(st), aux)
print("reads", nreads, "samples", ns, "depress ms", e0.elapsed_time(e1), "GB/s raw", 2 * ns / e0.elapsed_time(e1) / 1e6)
print("steps", aux[1], "rounds", aux[2], "rounds/step", aux[2] / max(1, aux[1]))
assert torch.equal(d_back[:ns], d_s[:ns])
