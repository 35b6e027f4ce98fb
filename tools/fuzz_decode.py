"""Corrupted-stream fuzz of the decoders (run on a GPU box): truncated, bit-flipped and
length-field-forged streams must come back as an error or as (wrong) samples, never as a CUDA
fault.  After every call a small valid round trip proves the context is still alive.
    python tools/fuzz_decode.py [cases per method] [seed]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from honours_b200 import press  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 150
SEED = int(sys.argv[2]) if len(sys.argv) > 2 else 1
press.init(0)
press.set_huffman_table()
rng = np.random.default_rng(SEED)
probe = (500 + np.cumsum(rng.integers(-20, 21, 3000))).astype(np.int16)


def alive(method):
    rc, st = press.press(method, probe)
    if rc != 0:
        return False
    rc, back = press.depress(method, st, probe.size)
    return rc == 0 and np.array_equal(back, probe)


def make(kind, n):
    d = rng.integers(-12, 13, n)
    if kind:
        idx = rng.random(n) < (0.01 if kind == 1 else 0.3)
        d[idx] = rng.integers(-900, 900, int(idx.sum()))
    return (500 + np.cumsum(d)).astype(np.int16)


bad = 0
for method in press.implemented_methods():
    errs = oks = 0
    for case in range(N):
        n = int(rng.integers(2, 20000))
        s = make(case % 3, n)
        rc, st = press.press(method, s)
        if rc != 0:
            continue
        b = bytearray(st)
        mode = case % 5
        if mode == 0 and len(b) > 1:                       # truncated
            b = b[: int(rng.integers(0, len(b)))]
        elif mode == 1:                                    # a few flipped bytes anywhere
            for _ in range(int(rng.integers(1, 6))):
                b[int(rng.integers(0, len(b)))] ^= int(rng.integers(1, 256))
        elif mode == 2:                                    # flipped bytes in the header
            for _ in range(int(rng.integers(1, 4))):
                b[int(rng.integers(0, min(len(b), 24)))] ^= int(rng.integers(1, 256))
        elif mode == 3 and len(b) >= 8:                    # forged 32-bit field near the front
            o = int(rng.integers(0, min(len(b) - 4, 20)))
            b[o:o + 4] = int(rng.choice([0xffffffff, 0x7fffffff, n, n + 1, 0x10000, 0])).to_bytes(4, "little")
        else:                                              # wrong sample count
            n = max(1, n + int(rng.integers(-3, 4)))
        try:
            rc, back = press.depress(method, bytes(b), n)
        except Exception as e:                             # the mirror raises on a dead context
            print(f"FAULT {method} case {case} mode {mode} n {n} len {len(b)}: {e}")
            sys.exit(1)
        errs += rc != 0
        oks += rc == 0
        if not alive(method):
            print(f"CONTEXT DEAD after {method} case {case} mode {mode} n {n} len {len(b)}: {press.last_error()}")
            sys.exit(1)
    print(f"{method}: {errs} rejected, {oks} decoded (to whatever the bytes say), context alive")
# ---- the batch path: a container of 40 reads with corrupted bytes / truncated tail
for method in press.implemented_methods():
    lens = rng.integers(2, 9000, 40)
    reads = [make(i % 3, int(n)) for i, n in enumerate(lens)]
    samples = np.concatenate(reads)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    rc, cont, oo = press.batch_press_host(method, samples, offsets)
    assert rc == 0
    rej = 0
    for case in range(max(4, N // 10)):
        c = cont.copy()
        for _ in range(int(rng.integers(1, 40))):
            c[int(rng.integers(0, c.size))] ^= int(rng.integers(1, 256))
        o2 = oo.copy()
        if case % 3 == 2:       # the last stream cut short
            o2[-1] = max(int(o2[-2]), int(o2[-1]) - int(rng.integers(1, 200)))
        try:
            press.batch_depress_host(method, c, o2, offsets)
        except press.PressGpuError:
            rej += 1
        if not alive(method):
            print(f"CONTEXT DEAD after batch {method} case {case}: {press.last_error()}")
            sys.exit(1)
    print(f"{method}: batch, {rej} containers rejected, context alive")
print("fuzz ok")
