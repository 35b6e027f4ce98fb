"""Records in, container out: the BLOW5 reader and the container file of
honours_b200/csrc/blow5_io.cu, through the C ABI."""
import hashlib
import json
import os
import struct
import zlib

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _write_blow5(path, reads, record_zlib):
    """A minimal BLOW5 v0.2.0 file with RAW int16 signals (signal-press 0), layout of
    press/sigtk/slow5lib/src/slow5.c:805-830 (header) and :3925-3960 (records)."""
    hdr = b"#slow5_version\t0.2.0\n#num_read_groups\t1\n"
    head = bytearray(64)
    head[:6] = b"BLOW5\x01"
    head[6:9] = bytes([0, 2, 0])
    head[9] = 1 if record_zlib else 0
    head[10:14] = struct.pack("<I", 1)
    head[14] = 0
    with open(path, "wb") as f:
        f.write(head)
        f.write(struct.pack("<I", len(hdr)))
        f.write(hdr)
        for rid, s in reads:
            rec = struct.pack("<H", len(rid)) + rid.encode() + struct.pack("<I", 0) + struct.pack("<4d", 2048, 3, 748, 4000)
            rec += struct.pack("<Q", s.size) + s.astype("<i2").tobytes()
            if record_zlib:
                rec = zlib.compress(rec)
            f.write(struct.pack("<Q", len(rec)) + rec)
        f.write(b"5WOLB")


@pytest.mark.parametrize("record_zlib", [False, True])
def test_blow5_reader_raw_signals(tmp_path, record_zlib):
    """raw-signal files need no GPU: the records' samples land in the batch layout, batch by batch"""
    from honours_b200 import press
    rng = np.random.default_rng(3)
    reads = [(f"read-{i:04d}", (480 + np.cumsum(rng.integers(-20, 21, int(n)))).astype(np.int16))
             for i, n in enumerate([5, 1, 3000, 70000, 12, 40000])]
    p = str(tmp_path / "t.blow5")
    _write_blow5(p, reads, record_zlib)
    samples, offsets, ids, sp = press.blow5_read_all(p, max_reads=4, max_samples=75000)     # several batches
    assert sp == 0 and ids == [r[0] for r in reads]
    assert np.diff(offsets.astype(np.int64)).tolist() == [r[1].size for r in reads]
    assert np.array_equal(samples, np.concatenate([r[1] for r in reads]))


def test_container_file_round_trip(tmp_path):
    from honours_b200 import press
    rng = np.random.default_rng(4)
    sizes = rng.integers(0, 5000, 40)
    oo = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    off = np.concatenate([[0], np.cumsum(rng.integers(1, 9000, 40))]).astype(np.uint64)
    cont = rng.integers(0, 256, int(oo[-1]), dtype=np.uint8)
    p = str(tmp_path / "c.prs")
    press.container_write(p, "ex_zd", off, oo, cont)
    m, off2, oo2, cont2 = press.container_read(p)
    assert m == "ex_zd" and np.array_equal(off, off2) and np.array_equal(oo, oo2) and np.array_equal(cont, cont2)
    open(p, "r+b").truncate(os.path.getsize(p) - 10)
    with pytest.raises(press.PressGpuError):
        press.container_read(p)


def test_blow5_fixture_header():
    """tests/golden/three-reads.blow5 (copied verbatim from the reference's data/: an input vector)
    keeps its signals svb-zd coded: reading it needs the GPU"""
    from honours_b200 import press
    L = press.lib()
    f = L.press_gpu_blow5_open(os.path.join(GOLD, "three-reads.blow5").encode())
    assert f and L.press_gpu_blow5_signal_press(f) == 1
    L.press_gpu_blow5_close(f)


@pytest.mark.gpu
@pytest.mark.parametrize("method", ["svb16_zd", "ex_zd", "shuff_vbe21_zd"])
def test_blow5_to_container_and_back(tmp_path, method):
    """config 1 end to end from the file: data/three-reads.blow5 -> records (frames decoded by the GPU)
    -> batch press -> container file -> read back -> depress == the signals; every stream has the size
    and sha256 the compiled reference produced (tests/golden/three_reads_ref.json)"""
    from honours_b200 import press
    press.init(0)
    press.set_huffman_table()
    samples, offsets, ids, sp = press.blow5_read_all(os.path.join(GOLD, "three-reads.blow5"))
    z = np.load(os.path.join(GOLD, "three_reads.npz"))
    assert sp == 1 and ids == [str(x) for x in z["ids"]]
    assert np.array_equal(offsets, z["offsets"]) and np.array_equal(samples, z["samples"])
    rc, cont, oo = press.batch_press_host(method, samples, offsets)
    assert rc == 0
    p = str(tmp_path / "three.prs")
    press.container_write(p, method, offsets, oo, cont)
    m, off2, oo2, cont2 = press.container_read(p)
    gold = json.load(open(os.path.join(GOLD, "three_reads_ref.json")))
    for r, rid in enumerate(ids):
        st = cont2[int(oo2[r]):int(oo2[r + 1])].tobytes()
        assert len(st) == gold[rid][method]["bytes"] and hashlib.sha256(st).hexdigest() == gold[rid][method]["sha256"]
    back = press.batch_depress_host(m, cont2, oo2, off2)
    assert m == method and np.array_equal(back, samples)
