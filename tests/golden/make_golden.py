#!/usr/bin/env python
"""Regenerate tests/golden/ from the reference itself (run in the build container).

Needs /root/reference and oracle/_ref/libpress_ref.so (``make -C oracle ref``).
Writes, all small enough to commit:

* three_reads.npz       -- the raw int16 signals of data/three-reads.blow5 (the
                           file stores them as svb-zd frames; they are decoded
                           with the REFERENCE's svb_zd_depress_16, never with our
                           code), their read ids, and the stored frames' FNV hashes
* three_reads_ref.json  -- per read, per method: stream size + FNV-1a-64 + sha256
                           of the stream the compiled reference produced
* small_streams.npz     -- complete reference streams for read 0 (7329 samples)
                           and for seeded edge-case reads (n = 1..70, exception
                           heavy, q-shifted), every method
* data files for the product: honours_b200/data/NA12878_zd.huffman (the
  committed table, a binary input named by north_star) and
  honours_b200/data/NA12878_zd_hist.json (the 256-bin one-byte-zd histogram of
  press/gen_huffman.c:11 -- the synthetic-data specification).
"""
import hashlib
import json
import os
import re
import struct
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import cpu_impl  # noqa: E402

REF = "/root/reference"


def read_blow5(path):
    """Minimal BLOW5 v0.2.0 reader (zlib records, svb-zd or raw signal).
    Layout per press/sigtk/slow5lib/src/slow5.c:805-830 (binary header) and
    :3947-3960 (record: the signal field holds the svb-zd frame when the
    header's signal-press byte is 1)."""
    b = open(path, "rb").read()
    assert b[:6] == b"BLOW5\x01"
    rec_press, = struct.unpack_from("<B", b, 9)
    sig_press, = struct.unpack_from("<B", b, 14)
    assert rec_press == 1, "zlib records expected"
    hdr_len, = struct.unpack_from("<I", b, 64)
    off = 68 + hdr_len
    out = []
    while b[off:off + 5] != b"5WOLB":
        size, = struct.unpack_from("<Q", b, off)
        off += 8
        rec = zlib.decompress(b[off:off + size])
        off += size
        idlen, = struct.unpack_from("<H", rec, 0)
        rid = rec[2:2 + idlen].decode()
        p = 2 + idlen + 4 + 4 * 8
        nfield, = struct.unpack_from("<Q", rec, p)
        p += 8
        if sig_press == 1:
            frame = rec[p:p + nfield]            # [n u32][classical svb over zd]
            out.append((rid, frame, None))
        else:
            out.append((rid, None, np.frombuffer(rec, "<i2", nfield, p).copy()))
    return out


def main():
    ref = cpu_impl.load("ref")
    table = open(f"{REF}/press/NA12878_zd.huffman", "rb").read()
    ref.set_huffman_table(table)

    os.makedirs(f"{ROOT}/honours_b200/data", exist_ok=True)
    open(f"{ROOT}/honours_b200/data/NA12878_zd.huffman", "wb").write(table)
    src = open(f"{REF}/press/gen_huffman.c").read()
    hist = [int(x) for x in re.search(r"\{([^}]*)\}", src).group(1).split(",")]
    assert len(hist) == 256 and sum(hist) == 56732772179
    json.dump({"source": "press/gen_huffman.c:11", "counts": hist},
              open(f"{ROOT}/benchdata/NA12878_zd_hist.json", "w"))

    # ---- three-reads.blow5 -------------------------------------------------
    reads = read_blow5(f"{REF}/data/three-reads.blow5")
    sigs, ids, frame_hash = [], [], []
    for rid, frame, raw in reads:
        if raw is None:
            n, = struct.unpack_from("<I", frame, 0)
            rc, raw = ref.depress("vbz_svb_zd", frame, n)
            assert rc == 0 and raw.size == n
            # the stored frame is a known-answer vector for the pre-zstd VBZ stream
            rc, again = ref.press("vbz_svb_zd", raw)
            assert rc == 0 and again == frame, "reference re-encode differs from file"
            frame_hash.append(f"{cpu_impl.fnv1a64(frame):016x}")
        sigs.append(raw)
        ids.append(rid)
    offsets = np.concatenate([[0], np.cumsum([s.size for s in sigs])]).astype(np.uint64)
    np.savez_compressed(f"{HERE}/three_reads.npz", samples=np.concatenate(sigs),
                        offsets=offsets, ids=np.array(ids), frame_fnv=np.array(frame_hash))

    summary = {}
    for rid, s in zip(ids, sigs):
        summary[rid] = {"n": int(s.size)}
        for m in cpu_impl.METHODS:
            rc, st = ref.press(m, s)
            assert rc == 0
            rc2, back = ref.depress(m, st, s.size)
            assert rc2 == 0 and np.array_equal(back, s), (m, rid)
            summary[rid][m] = {"bytes": len(st), "fnv1a64": f"{cpu_impl.fnv1a64(st):016x}",
                               "sha256": hashlib.sha256(st).hexdigest()}
    json.dump(summary, open(f"{HERE}/three_reads_ref.json", "w"), indent=1)

    # ---- small complete streams -------------------------------------------
    rng = np.random.default_rng(0x5EED0001)
    cases = {"read0": sigs[0]}
    for n in (1, 2, 3, 7, 8, 9, 15, 16, 17, 31, 32, 33, 63, 64, 65, 70):
        cases[f"walk_n{n}"] = (475 + np.cumsum(rng.integers(-20, 21, n))).astype(np.int16)
    cases["heavy_ex_300"] = rng.integers(-2000, 2000, 300).astype(np.int16)
    cases["one_ex_200"] = np.r_[np.full(100, 480), np.full(100, 900)].astype(np.int16)
    cases["two_ex_200"] = np.r_[np.full(50, 480), np.full(100, 900), np.full(50, 300)].astype(np.int16)
    cases["no_ex_500"] = (500 + rng.integers(-60, 61, 500)).astype(np.int16)
    cases["q2_400"] = ((120 + np.cumsum(rng.integers(-9, 10, 400))) * 4).astype(np.int16)
    cases["q5_cap_100"] = ((10 + rng.integers(0, 5, 100)) * 64).astype(np.int16)
    cases["wrap_300"] = rng.integers(-32768, 32768, 300).astype(np.int16)
    cases["big_gap_70000"] = np.r_[np.full(10, 100), np.full(69980, 500), np.full(10, 3000)].astype(np.int16)
    blobs = {}
    for name, s in cases.items():
        blobs[f"{name}/in"] = s
        for m in cpu_impl.METHODS:
            rc, st = ref.press(m, s)
            blobs[f"{name}/{m}/rc"] = np.int64(rc)
            blobs[f"{name}/{m}"] = np.frombuffer(st, dtype=np.uint8)
    np.savez_compressed(f"{HERE}/small_streams.npz", **blobs)
    print("golden written:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
