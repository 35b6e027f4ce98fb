"""GPU parity tests (run with -m gpu on a B200): every stream the CUDA path
writes must be byte-identical to the oracle's (== the reference's) stream for
the same read, and every decode must return the original samples.  All calls
go through the C ABI (include/press_gpu.h) via honours_b200.press."""
import json
import os

import numpy as np
import pytest

from conftest import narrow_reads, svb0124_stream

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
IMPLEMENTED = ["svb16_zd", "svb_zd", "vbz_svb16_zd", "vbz_svb_zd",
               "vbe21_zd", "vbsse21_zd", "ex_zd", "shuff_vbe21_zd", "shuff_vbsse21_zd", "vbbe21_zd", "shuff_vbbe21_zd",
               "svb0124_zd", "uint_zd"]
HUFF = ("shuff_vbe21_zd", "shuff_vbsse21_zd", "shuff_vbbe21_zd")
HDR16 = ("shuff_vbsse21_zd", "shuff_vbbe21_zd")    # header length kept in a uint16_t (press.c:4753, 4521)
EXFAM = IMPLEMENTED[4:11]


@pytest.fixture(scope="module")
def gpu():
    from honours_b200 import press
    press.init(0)
    press.set_huffman_table()
    assert press.implemented_methods() == IMPLEMENTED
    return press


@pytest.fixture(scope="module")
def three_reads():
    z = np.load(os.path.join(GOLD, "three_reads.npz"))
    return z["samples"], z["offsets"], [str(x) for x in z["ids"]]


def rand_read(rng, kind, n):
    if kind == "walk":
        return (475 + np.cumsum(rng.integers(-25, 26, n))).astype(np.int16)
    if kind == "spiky":
        d = rng.integers(-12, 13, n)
        idx = rng.random(n) < 0.01
        d[idx] = rng.integers(-900, 900, int(idx.sum()))
        return (500 + np.cumsum(d)).astype(np.int16)
    if kind == "q":
        return ((100 + np.cumsum(rng.integers(-6, 7, n))) << int(rng.integers(1, 7))).astype(np.int16)
    return rng.integers(-32768, 32768, n).astype(np.int16)


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_three_reads_per_read_dropin(gpu, oracle, checker, three_reads, method):
    """config 1: data/three-reads.blow5 through the press.h-shaped calls."""
    import hashlib
    samples, offsets, ids = three_reads
    gold = json.load(open(os.path.join(GOLD, "three_reads_ref.json")))
    for r, rid in enumerate(ids):
        s = samples[int(offsets[r]):int(offsets[r + 1])]
        rc, st = gpu.press(method, s)
        assert rc == 0
        assert len(st) == gold[rid][method]["bytes"]
        assert hashlib.sha256(st).hexdigest() == gold[rid][method]["sha256"]
        assert st == oracle.press(method, s)[1] == checker.press(method, s)[1]
        rc, back = gpu.depress(method, st, s.size)
        assert rc == 0 and np.array_equal(back, s)


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_edge_lengths_per_read(gpu, oracle, method):
    rng = np.random.default_rng(42)
    for n in [1, 2, 3, 7, 8, 9, 15, 16, 17, 63, 64, 65, 4095, 4096, 4097, 8191, 8192, 8193, 12289]:
        for kind in ("walk", "noise", "spiky", "q"):
            s = rand_read(rng, kind, n)
            rc, st = gpu.press(method, s)
            ro, so = oracle.press(method, s)
            # ex-zd gives up on streams longer than 2n + 1024 (ex_zd.c:411): same verdict expected
            assert (rc == 0) == (ro == 0), (method, kind, n, rc, ro)
            if ro != 0:
                continue
            assert st == so, (method, kind, n)
            ro, bo = oracle.depress(method, so, n)
            rc, back = gpu.depress(method, so, n)
            # (the reference cannot decode a Huffman stream with no symbols, huffman.c:696-711)
            assert (rc == 0) == (ro == 0), (method, kind, n, rc, ro)
            if ro == 0:
                assert np.array_equal(back, s), (method, kind, n)


def make_batch(rng, nreads, lo, hi, kinds=("walk", "spiky", "q", "noise")):
    lens = rng.integers(lo, hi, nreads)
    reads = [rand_read(rng, kinds[i % len(kinds)], int(n)) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    return np.concatenate(reads), offsets, reads


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_batch_host_matches_oracle(gpu, checker, method):
    """tight container: stream r at out_offsets[r], all byte-identical to the oracle;
    reads are back to back so most start at a misaligned address."""
    rng = np.random.default_rng(7)
    kinds = ("walk", "spiky", "q") if method in ("ex_zd",) + HUFF else ("walk", "spiky", "q", "noise")
    # (shuffman-vbsse21: coded exception lists must stay below 64 KB, press.c:4753)
    hi = 12000 if method in HDR16 else 30000
    samples, offsets, reads = make_batch(rng, 300, 2 if method in EXFAM else 1, hi, kinds=kinds)
    rc, cont, oo = gpu.batch_press_host(method, samples, offsets)
    assert rc == 0
    expect = [checker.press(method, s)[1] for s in reads]
    assert int(oo[0]) == 0
    assert np.diff(oo.astype(np.int64)).tolist() == [len(e) for e in expect]
    assert cont.tobytes() == b"".join(expect)
    back = gpu.batch_depress_host(method, cont, oo, offsets)
    assert np.array_equal(back, samples)


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_batch_with_empty_and_tiny_reads(gpu, oracle, method):
    rng = np.random.default_rng(9)
    lens = [0, 2 if method in HUFF else 1, 0, 5, 4096, 0, 4097, 8, 0]
    reads = [rand_read(rng, "spiky", n) for n in lens]
    samples = np.concatenate(reads) if reads else np.zeros(0, np.int16)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    rc, cont, oo = gpu.batch_press_host(method, samples, offsets)
    assert rc == 0
    expect = [oracle.press(method, s)[1] if s.size else (b"\0\0\0\0" if method.startswith("vbz") else b"")
              for s in reads]
    assert cont.tobytes() == b"".join(expect)
    back = gpu.batch_depress_host(method, cont, oo, offsets)
    assert np.array_equal(back, samples)


def test_long_read_many_tiles(gpu, oracle):
    """one 3.3 M-sample read: > 800 tiles chained by look-back inside one read"""
    rng = np.random.default_rng(11)
    s = rand_read(rng, "spiky", 3_300_001)
    for method in ("svb16_zd", "vbz_svb_zd", "vbe21_zd", "vbsse21_zd", "ex_zd", "shuff_vbe21_zd", "vbbe21_zd",
                   "svb0124_zd", "uint_zd"):
        rc, st = gpu.press(method, s)
        assert rc == 0 and st == oracle.press(method, s)[1]
        rc, back = gpu.depress(method, st, s.size)
        assert rc == 0 and np.array_equal(back, s)


@pytest.mark.parametrize("method", ["svb0124_zd", "uint_zd", "svb16_zd", "shuff_vbe21_zd"])
def test_narrow_and_sparse_values(gpu, checker, method):
    rng = np.random.default_rng(77)
    reads = narrow_reads(rng)
    if method.startswith("shuff"):     # (a stream without payload symbols cannot be read back, huffman.c:706: tested elsewhere)
        reads = [r for r in reads if r.size > 2]
    for s in reads:     # per read (the press.h-shaped calls)
        rc, st = gpu.press(method, s)
        rc0, want = checker.press(method, s)
        assert rc == rc0 and st == want, (method, s.size)
        if rc == 0:
            rc, back = gpu.depress(method, st, s.size)
            assert rc == 0 and np.array_equal(back, s), (method, s.size)
    samples = np.concatenate(reads)     # and as one batch (streams of all phases next to each other)
    offsets = np.concatenate([[0], np.cumsum([r.size for r in reads])]).astype(np.uint64)
    rc, cont, oo = gpu.batch_press_host(method, samples, offsets)
    expect = [checker.press(method, s) for s in reads]
    if all(e[0] == 0 for e in expect):
        assert rc == 0 and cont.tobytes() == b"".join(e[1] for e in expect)
        assert np.array_equal(gpu.batch_depress_host(method, cont, oo, offsets), samples)


@pytest.mark.parametrize("n,p4", [(5000, 0.002), (9000, 0.02), (12000, 0.55), (4099, 0.58)])
def test_svb0124_decodes_four_byte_values(gpu, checker, n, p4):
    """svb0124-zd decode of values wider than 16 bits: a few among many (the per-thread 32-bit path) and
    tiles that hold more bytes than the shared-memory stage (read from the container byte by byte)"""
    rng = np.random.default_rng(n)
    st = svb0124_stream(rng, n, p4)
    rc0, want = checker.depress("svb0124_zd", st, n)
    rc, got = gpu.depress("svb0124_zd", st, n)
    assert rc == rc0 == 0 and np.array_equal(got, want)


@pytest.mark.parametrize("method", ["svb16_zd", "vbe21_zd", "vbsse21_zd", "ex_zd", "shuff_vbe21_zd", "vbbe21_zd",
                                    "svb0124_zd", "uint_zd"])
def test_device_batch_api_torch_pointers(gpu, checker, method):
    """the measured path: device-resident in/out, raw pointers through the C ABI"""
    import torch
    rng = np.random.default_rng(5)
    samples, offsets, reads = make_batch(rng, 600, 2000, 60000, kinds=("walk", "spiky", "q"))
    dev = torch.device("cuda:0")
    d_s = torch.from_numpy(samples).to(dev)
    d_off = torch.from_numpy(offsets.astype(np.int64)).to(dev)
    nreads, ns = len(reads), samples.size
    cap = gpu.batch_bound(method, nreads, ns)
    gpu.set_exception_capacity(1.0)     # the "q" reads are mostly exceptions for the methods without qts
    d_out = torch.empty(cap, dtype=torch.uint8, device=dev)
    d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
    sb = gpu.batch_scratch_bytes(method, nreads, ns)
    gpu.set_exception_capacity(1.0 / 64)
    d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    gpu.batch_press(method, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap,
                    d_oo.data_ptr(), d_scr.data_ptr(), sb, stream)
    assert gpu.batch_status(d_scr.data_ptr(), stream) == 0
    oo = d_oo.cpu().numpy()
    total = int(oo[-1])
    expect = b"".join(checker.press(method, s)[1] for s in reads)
    assert total == len(expect)
    assert d_out[:total].cpu().numpy().tobytes() == expect
    d_back = torch.zeros(ns, dtype=torch.int16, device=dev)
    gpu.batch_depress(method, d_out.data_ptr(), d_oo.data_ptr(), total, d_off.data_ptr(), nreads, ns,
                      d_back.data_ptr(), d_scr.data_ptr(), sb, stream)
    assert gpu.batch_status(d_scr.data_ptr(), stream) == 0
    assert torch.equal(d_back, d_s)


def test_capacity_error_is_reported(gpu):
    rng = np.random.default_rng(3)
    s = rand_read(rng, "noise", 5000)
    import ctypes as C
    out = np.zeros(100, dtype=np.uint8)
    oo = np.zeros(2, dtype=np.uint64)
    offs = np.array([0, s.size], dtype=np.uint64)
    rc = gpu.lib().press_gpu_batch_press_host(0, s.ctypes.data, offs.ctypes.data, 1,
                                              out.ctypes.data, 100, oo.ctypes.data)
    assert rc != 0


def test_ex_zd_scratch_cap_failure_mirrors_reference(gpu, oracle):
    """ex_zd.c:411: a stream longer than 2n + 1024 makes the reference return -1"""
    rng = np.random.default_rng(13)
    s = rand_read(rng, "noise", 20000)
    assert oracle.press("ex_zd", s)[0] != 0
    assert gpu.press("ex_zd", s)[0] != 0
    # inside a batch the read gets an empty stream and the call reports 1
    ok = rand_read(rng, "walk", 5000)
    samples = np.concatenate([ok, s, ok])
    offsets = np.array([0, 5000, 25000, 30000], dtype=np.uint64)
    rc, cont, oo = gpu.batch_press_host("ex_zd", samples, offsets)
    e = oracle.press("ex_zd", ok)[1]
    assert rc == 1 and cont.tobytes() == e + e and int(oo[1]) == int(oo[2])


@pytest.mark.parametrize("method", HDR16)
def test_shuffman_header_limit_mirrors_reference(gpu, oracle, method):
    """press.c:4753 / 4521 keep the header length in a uint16_t: once the coded exception lists pass
    65535 bytes the reference writes a stream it cannot read back; both sides refuse the read"""
    rng = np.random.default_rng(14)
    s = rand_read(rng, "noise", 40000)          # ~40000 exceptions, > 130 KB of coded lists
    assert oracle.press(method, s)[0] != 0
    assert gpu.press(method, s)[0] != 0
    small = rand_read(rng, "noise", 15000)      # ~50 KB of lists: still representable
    ro, so = oracle.press(method, small)
    rc, st = gpu.press(method, small)
    assert ro == rc == 0 and st == so
    ok = rand_read(rng, "walk", 5000)
    samples = np.concatenate([ok, s, ok])
    offsets = np.array([0, 5000, 45000, 50000], dtype=np.uint64)
    rc, cont, oo = gpu.batch_press_host(method, samples, offsets)
    e = oracle.press(method, ok)[1]
    assert rc == 1 and cont.tobytes() == e + e and int(oo[1]) == int(oo[2])


@pytest.mark.parametrize("method", EXFAM)
@pytest.mark.parametrize("p_ex", [0.0, 1e-3, 0.1, 0.5])
def test_exception_rate_sweep(gpu, checker, method, p_ex):
    """config 3: exception-rate sweep on synthetic NA12878-like reads (host generator)"""
    from benchdata import workload
    lengths, x0, shift = workload.read_table(40, q_fraction=0.25)
    lengths = np.minimum(lengths, 12000 if method in HDR16 else 30000)    # (press.c:4753)
    offsets = workload.offsets_of(lengths)
    samples = workload.synth_host(offsets, x0, shift, p_ex=p_ex)
    reads = [samples[int(offsets[r]):int(offsets[r + 1])] for r in range(len(lengths))]
    rc, cont, oo = gpu.batch_press_host(method, samples, offsets)
    assert rc == 0
    expect = [checker.press(method, s)[1] for s in reads]
    assert np.diff(oo.astype(np.int64)).tolist() == [len(e) for e in expect]
    assert cont.tobytes() == b"".join(expect)
    back = gpu.batch_depress_host(method, cont, oo, offsets)
    assert np.array_equal(back, samples)


def test_huffman_decodes_oracle_streams_without_sidecar(gpu, oracle):
    """config 4: streams written by the CPU reference path carry no sub-block offsets;
    the self-synchronising decoder must recover them exactly"""
    rng = np.random.default_rng(17)
    for n in (2, 30, 1000, 4097, 70001, 300001):
        s = rand_read(rng, "walk" if n % 2 else "spiky", n)
        ro, so = oracle.press("shuff_vbe21_zd", s)
        assert ro == 0
        rc, back = gpu.depress("shuff_vbe21_zd", so, n)
        assert rc == 0 and np.array_equal(back, s), n


def _device_roundtrip(gpu, method, samples, offsets, sidecar=True, corrupt=None):
    """device batch press (+ side-car) and depress -> (container bytes, out_offsets, decoded, status)"""
    import torch
    dev = torch.device("cuda:0")
    nreads, ns = offsets.size - 1, int(offsets[-1])
    d_s = torch.from_numpy(samples).to(dev)
    d_off = torch.from_numpy(offsets.astype(np.int64)).to(dev)
    cap = gpu.batch_bound(method, nreads, ns)
    gpu.set_exception_capacity(1.0)
    sb = gpu.batch_scratch_bytes(method, nreads, ns)
    gpu.set_exception_capacity(1.0 / 64)
    nsc = gpu.sidecar_bytes(method, nreads, ns) if sidecar else 0
    d_out = torch.empty(cap + 64, dtype=torch.uint8, device=dev)
    d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
    d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    d_sc = torch.zeros(nsc + 64, dtype=torch.uint8, device=dev) if nsc else None
    st = torch.cuda.current_stream().cuda_stream
    gpu.batch_press(method, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap, d_oo.data_ptr(),
                    d_scr.data_ptr(), sb, st, sidecar=d_sc.data_ptr() if nsc else 0, sidecar_bytes=nsc)
    assert gpu.batch_status(d_scr.data_ptr(), st) == 0
    oo = d_oo.cpu().numpy().astype(np.uint64)
    total = int(oo[-1])
    if corrupt is not None and nsc:
        corrupt(d_sc)
    d_back = torch.full((ns + 64,), -1, dtype=torch.int16, device=dev)
    gpu.batch_depress(method, d_out.data_ptr(), d_oo.data_ptr(), total, d_off.data_ptr(), nreads, ns,
                      d_back.data_ptr(), d_scr.data_ptr(), sb, st, sidecar=d_sc.data_ptr() if nsc else 0,
                      sidecar_bytes=nsc)
    status = gpu.batch_status(d_scr.data_ptr(), st)
    assert bool((d_back[ns:] == -1).all())
    return d_out[:total].cpu().numpy().tobytes(), oo, d_back[:ns].cpu().numpy(), status


@pytest.mark.parametrize("method", HUFF)
def test_huffman_sidecar_decode(gpu, checker, method):
    """config 4, our own streams: press leaves the side-car (bit offset of every 32nd payload symbol)
    OUTSIDE the container, whose bytes stay the reference's; depress with it (fused bits -> samples
    kernel) returns the input; the same container decodes without it as well (self-synchronising)."""
    rng = np.random.default_rng(31)
    lens = [2, 3, 33, 34, 35, 65, 4096, 4097, 8191, 8192, 8193, 8225, 8226, 12289, 40000, 70001, 200000, 1, 0, 9]
    kinds = ("walk", "spiky", "q", "walk")
    hi = 12000 if method in HDR16 else 10 ** 9
    reads = [rand_read(rng, kinds[i % 4], min(n, hi) if kinds[i % 4] != "walk" else n) for i, n in enumerate(lens)]
    reads = [r if r.size != 1 else rand_read(rng, "walk", 2) for r in reads]       # (no 1-sample Huffman streams)
    samples = np.concatenate(reads)
    offsets = np.concatenate([[0], np.cumsum([r.size for r in reads])]).astype(np.uint64)
    cont, oo, back, status = _device_roundtrip(gpu, method, samples, offsets, sidecar=True)
    assert status == 0
    assert cont == b"".join(checker.press(method, r)[1] if r.size else b"" for r in reads)
    assert np.array_equal(back, samples)
    cont2, _, back2, status2 = _device_roundtrip(gpu, method, samples, offsets, sidecar=False)
    assert status2 == 0 and cont2 == cont and np.array_equal(back2, samples)


def test_huffman_sidecar_mismatch_is_reported(gpu):
    """a side-car that does not belong to the stream: status bit 1, no fault"""
    rng = np.random.default_rng(32)
    reads = [rand_read(rng, "walk", 50000), rand_read(rng, "spiky", 30000)]
    samples = np.concatenate(reads)
    offsets = np.array([0, 50000, 80000], dtype=np.uint64)

    def corrupt(d_sc):
        d_sc[64:4096] = 0x5a

    _, _, _, status = _device_roundtrip(gpu, "shuff_vbe21_zd", samples, offsets, corrupt=corrupt)
    assert status & 2
    _, _, back, status = _device_roundtrip(gpu, "shuff_vbe21_zd", samples, offsets)
    assert status == 0 and np.array_equal(back, samples)


def test_huffman_sidecar_long_codes_take_the_global_path(gpu, oracle):
    """a step whose codes average more than 12 bits does not fit the bit stage: decoded from global memory"""
    freq = [2.0 ** -min(k, 23) + 1e-9 * (256 - k) for k in range(256)]
    blob, lens = _huffman_table_blob(freq)
    sym = max(range(256), key=lambda s: (lens[s], s))
    delta = (sym >> 1) ^ -(sym & 1)
    na12878 = open(os.path.join(os.path.dirname(GOLD), "..", "honours_b200", "data", "NA12878_zd.huffman"), "rb").read()
    rng = np.random.default_rng(33)
    try:
        oracle.set_huffman_table(blob)
        gpu.set_huffman_table(blob)
        reads = [(np.arange(30000, dtype=np.int64) * delta).astype(np.int16), rand_read(rng, "walk", 20000)]
        samples = np.concatenate(reads)
        offsets = np.array([0, 30000, 50000], dtype=np.uint64)
        cont, oo, back, status = _device_roundtrip(gpu, "shuff_vbe21_zd", samples, offsets)
        assert status == 0 and np.array_equal(back, samples)
        assert cont == b"".join(oracle.press("shuff_vbe21_zd", r)[1] for r in reads)
    finally:
        oracle.set_huffman_table(na12878)
        gpu.set_huffman_table(na12878)


def _huffman_table_blob(freq):
    """A code-table file image (huffman.c:440-480 layout) of the Huffman code of `freq`:
    [count u32 BE][data bytes u32 BE] then per symbol [sym][nbits][code bytes, bit k of the
    code at bit k % 8 of byte k / 8]."""
    import heapq
    import struct
    heap = [(float(f), i, (i,)) for i, f in enumerate(freq)]
    heapq.heapify(heap)
    lens = [0] * 256
    tick = 256
    while len(heap) > 1:
        fa, _, sa = heapq.heappop(heap)
        fb, _, sb = heapq.heappop(heap)
        for s in sa + sb:
            lens[s] += 1
        heapq.heappush(heap, (fa + fb, tick, sa + sb))
        tick += 1
    assert max(lens) <= 24 and min(lens) >= 1
    blob = struct.pack(">II", 256, 0)
    code = 0
    prev = 0
    for ln, sym in sorted((lens[s], s) for s in range(256)):    # canonical codes, first bit = MSB
        code <<= ln - prev
        prev = ln
        bits = [(code >> (ln - 1 - k)) & 1 for k in range(ln)]
        raw = bytearray((ln + 7) // 8)
        for k, b in enumerate(bits):
            raw[k // 8] |= b << (k % 8)
        blob += bytes([sym, ln]) + bytes(raw)
        code += 1
    return blob, lens


@pytest.mark.parametrize("shape", ["short_codes", "long_codes"])
def test_huffman_custom_tables(gpu, oracle, shape):
    """other code tables than NA12878_zd.huffman: one whose shortest code has 1 bit (more
    symbols per 4 KB step than the symbol stage holds: the generic decoder) and one with
    >= 4-bit codes and a long tail beyond the 12-bit first-level table"""
    if shape == "short_codes":
        freq = [2.0 ** -min(k, 22) + 1e-9 * (256 - k) for k in range(256)]
    else:
        freq = [1.0] * 16 + [2.0 ** -(3 + k / 20) for k in range(240)]
    blob, lens = _huffman_table_blob(freq)
    assert (min(lens) < 4) == (shape == "short_codes") and max(lens) > 12
    na12878 = open(os.path.join(os.path.dirname(GOLD), "..", "honours_b200", "data", "NA12878_zd.huffman"), "rb").read()
    rng = np.random.default_rng(99)
    try:
        oracle.set_huffman_table(blob)
        gpu.set_huffman_table(blob)
        reads = [np.full(40000, 7, np.int16), rand_read(rng, "walk", 50001), rand_read(rng, "spiky", 9000),
                 (500 + np.cumsum(rng.integers(-1, 2, 120000))).astype(np.int16), rand_read(rng, "noise", 3000)]
        for s in reads:
            ro, so = oracle.press("shuff_vbe21_zd", s)
            rc, st = gpu.press("shuff_vbe21_zd", s)
            assert rc == ro == 0 and st == so
            rc, back = gpu.depress("shuff_vbe21_zd", so, s.size)
            assert rc == 0 and np.array_equal(back, s)
        lengths = [r.size for r in reads]
        offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(np.uint64)
        samples = np.concatenate(reads)
        rc, cont, oo = gpu.batch_press_host("shuff_vbe21_zd", samples, offsets)
        assert rc == 0
        assert cont.tobytes() == b"".join(oracle.press("shuff_vbe21_zd", s)[1] for s in reads)
        assert np.array_equal(gpu.batch_depress_host("shuff_vbe21_zd", cont, oo, offsets), samples)
    finally:
        oracle.set_huffman_table(na12878)
        gpu.set_huffman_table(na12878)


def test_huffman_tile_of_longest_24bit_codes(gpu, oracle):
    """a table whose longest code has 24 bits (the longest accepted) and reads made of
    that symbol alone: 4096-sample tiles of 24-bit codes must fit the encoder's stage."""
    freq = [2.0 ** -min(k, 23) + 1e-9 * (256 - k) for k in range(256)]
    blob, lens = _huffman_table_blob(freq)
    assert max(lens) == 24
    sym = max(range(256), key=lambda s: (lens[s], s))       # a one-byte zd value with a 24-bit code
    delta = (sym >> 1) ^ -(sym & 1)                         # unzigzag
    na12878 = open(os.path.join(os.path.dirname(GOLD), "..", "honours_b200", "data", "NA12878_zd.huffman"), "rb").read()
    try:
        oracle.set_huffman_table(blob)
        gpu.set_huffman_table(blob)
        for n in (4097, 20000):
            s = (np.arange(n, dtype=np.int64) * delta).astype(np.int16)
            ro, so = oracle.press("shuff_vbe21_zd", s)
            rc, st = gpu.press("shuff_vbe21_zd", s)
            assert ro == rc == 0 and st == so and len(st) > 3 * (n - 1) - 8
            rc, back = gpu.depress("shuff_vbe21_zd", so, n)
            assert rc == 0 and np.array_equal(back, s)
    finally:
        oracle.set_huffman_table(na12878)
        gpu.set_huffman_table(na12878)


def test_ex_zd_q_read_with_mostly_exceptions(gpu, oracle):
    """an ex-zd read with q > 0 whose shifted signal is mostly exceptions: both passes of
    the count sweep capture into the pool; the reference makes one pass (ex_zd.c:436-447) and succeeds."""
    rng = np.random.default_rng(21)
    for n in (1000, 30000):
        s = (np.cumsum(rng.choice([-200, 200], n)) << 2).astype(np.int16)
        ro, so = oracle.press("ex_zd", s)
        rc, st = gpu.press("ex_zd", s)
        assert (rc == 0) == (ro == 0)
        if ro == 0:
            assert st == so
            rc, back = gpu.depress("ex_zd", so, n)
            assert rc == 0 and np.array_equal(back, s)
    # inside a batch, next to ordinary reads
    reads = [rand_read(rng, "walk", 9000), (np.cumsum(rng.choice([-150, 150], 20000)) << 3).astype(np.int16),
             rand_read(rng, "q", 7000)]
    offsets = np.concatenate([[0], np.cumsum([r.size for r in reads])]).astype(np.uint64)
    rc, cont, oo = gpu.batch_press_host("ex_zd", np.concatenate(reads), offsets)
    assert rc == 0 and cont.tobytes() == b"".join(oracle.press("ex_zd", r)[1] for r in reads)
    assert np.array_equal(gpu.batch_depress_host("ex_zd", cont, oo, offsets), np.concatenate(reads))


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_bench_scale_chunk_properties(gpu, checker, method):
    """the bench's workload at chunk scale (9000 NA12878-like reads, ~1 G samples, generated on the
    device): decode(encode(x)) == x for every sample, the container is tight, and EVERY read's
    stream is byte-identical to the reference's: per-read length and FNV-1a-64 of the container
    slice against the same fingerprint of the stream the CPU checker writes for that read
    (oracle/_ref, the reference compiled in place, when it travelled to this box)."""
    import torch
    from benchdata import workload
    nreads = 9000
    dev = torch.device("cuda:0")
    lengths, x0, shift = workload.read_table(nreads, q_fraction=0.25 if method == "ex_zd" else 0.0)
    off = workload.offsets_of(lengths)
    ns = int(off[-1])
    d_off = torch.from_numpy(off.astype(np.int64)).to(dev)
    d_x0 = torch.from_numpy(x0).to(dev)
    d_sh = torch.from_numpy(shift).to(dev)
    d_s = torch.empty(ns + 64, dtype=torch.int16, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    workload.synth_device(d_s.data_ptr(), d_off.data_ptr(), nreads, d_x0.data_ptr(), d_sh.data_ptr(), stream=st)
    cap = gpu.batch_bound(method, nreads, ns)
    sb = gpu.batch_scratch_bytes(method, nreads, ns)
    d_out = torch.empty(cap, dtype=torch.uint8, device=dev)
    d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
    d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    nsc = gpu.sidecar_bytes(method, nreads, ns)     # (Huffman methods: decode through the fused side-car kernel)
    d_sc = torch.empty(nsc + 64, dtype=torch.uint8, device=dev) if nsc else None
    gpu.batch_press(method, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap,
                    d_oo.data_ptr(), d_scr.data_ptr(), sb, st, sidecar=d_sc.data_ptr() if nsc else 0, sidecar_bytes=nsc)
    assert gpu.batch_status(d_scr.data_ptr(), st) == 0
    oo = d_oo.cpu().numpy().astype(np.uint64)
    assert oo[0] == 0 and np.all(np.diff(oo.astype(np.int64)) > 0) and int(oo[-1]) <= cap
    samples = d_s[:ns].cpu().numpy()
    cont = d_out[: int(oo[-1])].cpu().numpy()
    want_h, want_sz = checker.hash_batch(method, samples, off)
    assert np.array_equal(np.diff(oo.astype(np.int64)).astype(np.uint64), want_sz), method
    got_h = checker.hash_slices(cont, oo)
    bad = np.nonzero(got_h != want_h)[0]
    assert bad.size == 0, (method, checker.kind, bad[:8].tolist())
    del samples, cont
    d_back = torch.full((ns + 64,), -1, dtype=torch.int16, device=dev)
    gpu.batch_depress(method, d_out.data_ptr(), d_oo.data_ptr(), int(oo[-1]), d_off.data_ptr(), nreads, ns,
                      d_back.data_ptr(), d_scr.data_ptr(), sb, st, sidecar=d_sc.data_ptr() if nsc else 0, sidecar_bytes=nsc)
    assert gpu.batch_status(d_scr.data_ptr(), st) == 0
    assert torch.equal(d_back[:ns], d_s[:ns])
    assert bool((d_back[ns:] == -1).all())      # nothing written past the last sample
    if nsc:     # and once more without the side-car: the self-synchronising decoder on the same container
        d_back.fill_(-1)
        gpu.batch_depress(method, d_out.data_ptr(), d_oo.data_ptr(), int(oo[-1]), d_off.data_ptr(), nreads, ns,
                          d_back.data_ptr(), d_scr.data_ptr(), sb, st)
        assert gpu.batch_status(d_scr.data_ptr(), st) == 0
        assert torch.equal(d_back[:ns], d_s[:ns])
    del d_s, d_out, d_back, d_scr, d_sc
    torch.cuda.empty_cache()


def test_corrupted_streams_do_not_fault():
    """truncated, bit-flipped and length-forged streams through every per-read depress: an error
    or wrong samples, never a crash or a CUDA fault (own process: a fault would poison the context
    of this session).  tools/fuzz_decode.py checks after every call that the context still works."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "fuzz_decode.py"), "40", "7"],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0 and "fuzz ok" in r.stdout, (r.returncode, r.stdout[-800:], r.stderr[-800:])
