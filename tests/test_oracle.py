"""CPU tests: pin the oracle (our restatement) to the reference.

1. against tests/golden/ (streams produced by the compiled reference, committed);
2. against oracle/_ref (the reference compiled in place) on seeded random reads,
   when that library exists (it does in the build container and travels to the
   GPU box; /root/reference itself is never read at test time).
"""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import cpu_impl

GOLD = os.path.join(os.path.dirname(__file__), "golden")
METHODS = list(cpu_impl.METHODS)


@pytest.fixture(scope="module")
def three_reads():
    z = np.load(os.path.join(GOLD, "three_reads.npz"))
    return z["samples"], z["offsets"], [str(x) for x in z["ids"]]


def test_three_reads_shape(three_reads):
    samples, offsets, ids = three_reads
    assert np.diff(offsets).tolist() == [7329, 155185, 95350]
    assert samples.dtype == np.int16 and samples.size == 257864


@pytest.mark.parametrize("method", METHODS)
def test_oracle_matches_reference_golden_three_reads(oracle, three_reads, method):
    samples, offsets, ids = three_reads
    gold = json.load(open(os.path.join(GOLD, "three_reads_ref.json")))
    for r, rid in enumerate(ids):
        s = samples[int(offsets[r]):int(offsets[r + 1])]
        rc, st = oracle.press(method, s)
        assert rc == 0
        g = gold[rid][method]
        assert len(st) == g["bytes"]
        assert hashlib.sha256(st).hexdigest() == g["sha256"]
        rc, back = oracle.depress(method, st, s.size)
        assert rc == 0 and np.array_equal(back, s)


def test_blow5_frames_are_vbz_svb_zd(oracle, three_reads):
    """The .blow5 fixture stores each signal as [n u32][svb32 over zd]: a
    known-answer vector written by slow5lib, not by press.c."""
    samples, offsets, ids = three_reads
    z = np.load(os.path.join(GOLD, "three_reads.npz"))
    for r in range(3):
        s = samples[int(offsets[r]):int(offsets[r + 1])]
        rc, st = oracle.press("vbz_svb_zd", s)
        assert rc == 0 and f"{cpu_impl.fnv1a64(st):016x}" == str(z["frame_fnv"][r])


def _small_cases():
    z = np.load(os.path.join(GOLD, "small_streams.npz"))
    names = sorted({k.split("/")[0] for k in z.files})
    return z, names


@pytest.mark.parametrize("method", METHODS)
def test_oracle_matches_reference_small_streams(oracle, method):
    z, names = _small_cases()
    for name in names:
        s = z[f"{name}/in"]
        ref_rc = int(z[f"{name}/{method}/rc"])
        ref_st = z[f"{name}/{method}"].tobytes()
        rc, st = oracle.press(method, s)
        assert (rc == 0) == (ref_rc == 0), (name, rc, ref_rc)
        if rc != 0:
            continue
        assert st == ref_st, (name, method)
        if method in HUFF and len(st) - _huff_header_len(method, st) <= 4:
            assert oracle.depress(method, st, s.size)[0] != 0   # huffman.c:706 memread quirk
            continue
        rc, back = oracle.depress(method, st, s.size)
        assert rc == 0 and np.array_equal(back, s), (name, method)


HUFF = ("shuff_vbe21_zd", "shuff_vbsse21_zd", "shuff_vbbe21_zd")


def _huff_header_len(method, st):
    """bytes before [count u32 BE]: zd0, nex and the exception lists (press.c:4437, 4766-4775)"""
    nex = int.from_bytes(st[2:6], "little")
    if method == "shuff_vbe21_zd":
        return 6 + 6 * nex
    if nex == 1:
        return 12
    if nex == 0:
        return 6
    lp = int.from_bytes(st[6:10], "little")
    lv = int.from_bytes(st[10 + lp:14 + lp], "little")
    return 14 + lp + lv


def _random_read(rng, kind, n):
    if kind == "walk":
        return (475 + np.cumsum(rng.integers(-25, 26, n))).astype(np.int16)
    if kind == "spiky":
        d = rng.integers(-12, 13, n)
        idx = rng.random(n) < 0.01
        d[idx] = rng.integers(-900, 900, int(idx.sum()))
        return (500 + np.cumsum(d)).astype(np.int16)
    if kind == "q":
        return ((100 + np.cumsum(rng.integers(-6, 7, n))) << int(rng.integers(1, 7))).astype(np.int16)
    return rng.integers(-32768, 32768, n).astype(np.int16)       # "noise": mostly exceptions


@pytest.mark.parametrize("method", METHODS)
def test_oracle_matches_compiled_reference_random(oracle, ref, method):
    rng = np.random.default_rng(1234)
    for trial in range(60):
        kind = ["walk", "spiky", "q", "noise"][trial % 4]
        n = int(rng.integers(1, 6000)) if trial % 5 else int(rng.integers(1, 40))
        s = _random_read(rng, kind, n)
        rc_r, st_r = ref.press(method, s)
        rc_o, st_o = oracle.press(method, s)
        assert (rc_r == 0) == (rc_o == 0), (kind, n)
        if rc_r != 0:
            continue
        assert st_o == st_r, (method, kind, n)
        if method in HUFF and len(st_o) - _huff_header_len(method, st_o) <= 4:
            continue
        rc2, b = oracle.depress(method, st_r, n)
        assert rc2 == 0 and np.array_equal(b, s), (method, kind, n)
        if method in HUFF:  # press.c:4472 / 4821 size the scratch 2n bytes: the reference overflows it
            nex = int.from_bytes(st_o[2:6], "little")
            if _huff_header_len(method, st_o) - 2 + (n - 1 - nex) > 2 * n:
                continue
        rc, a = ref.depress(method, st_o, n)
        assert rc == 0 and np.array_equal(a, s), (method, kind, n)


def test_scalar_twin_nosimd_equals_simd(oracle, ref):
    """svb12_zd_nosimd_* (press.c:1647-1674, scalar streamvbyte_encode_12)
    writes the bytes svb12_zd_* (SSE svb16) writes, and both equal the restatement; each decodes
    the other's stream."""
    rng = np.random.default_rng(4321)
    for trial in range(80):
        kind = ["walk", "spiky", "q", "noise"][trial % 4]
        n = int(rng.integers(1, 9000)) if trial % 5 else int(rng.integers(1, 40))
        s = _random_read(rng, kind, n)
        scalar = ref.nosimd_press(s)
        rc, simd = ref.press("svb16_zd", s)
        assert rc == 0 and scalar == simd == oracle.press("svb16_zd", s)[1], (kind, n)
        assert np.array_equal(ref.nosimd_depress(simd, n), s)
        rc, back = ref.depress("svb16_zd", scalar, n)
        assert rc == 0 and np.array_equal(back, s)


def test_narrow_and_sparse_values_pinned(oracle, ref):
    """constant, staircase, rare-jump and sawtooth signals (uint-zd widths 0 .. 4, mostly-zero svb0124
    values, Huffman runs of one symbol): the restatement writes the compiled reference's bytes for every
    method, and each decodes the other's stream"""
    from conftest import narrow_reads
    reads = narrow_reads(np.random.default_rng(77))
    for method in cpu_impl.METHODS:
        for s in reads:
            rc0, want = ref.press(method, s)
            rc, got = oracle.press(method, s)
            assert rc == rc0 and got == want, (method, s.size)
            if rc == 0 and not (method.startswith("shuff") and s.size <= 2):
                a = oracle.depress(method, want, s.size)
                b = ref.depress(method, got, s.size)
                assert a[0] == b[0] == 0 and np.array_equal(a[1], s) and np.array_equal(b[1], s), (method, s.size)


def test_svb0124_four_byte_values_pinned(oracle, ref):
    """svb0124-zd decode of hand-made streams with 4-byte values (the reference's encoder never writes
    one for a 16-bit signal; its decoder takes them and keeps their low 16 bits, trans.c:272-282 ->
    unzigzag_one_16): the restatement follows the compiled reference"""
    from conftest import svb0124_stream
    for n, p4 in [(1, 0.5), (7, 0.4), (5000, 0.002), (9000, 0.02), (12000, 0.55), (4099, 0.58)]:
        st = svb0124_stream(np.random.default_rng(n), n, p4)
        rc0, want = ref.depress("svb0124_zd", st, n)
        rc, got = oracle.depress("svb0124_zd", st, n)
        assert rc == rc0 == 0 and np.array_equal(got, want), (n, p4)


def test_hash_batch_matches_per_read_hashes(oracle, ref, three_reads):
    """the multi-threaded fingerprint used by the bench-scale GPU parity test"""
    samples, offsets, _ = three_reads
    for impl in (oracle, ref):
        for m in ("svb16_zd", "ex_zd", "shuff_vbe21_zd"):
            h, sz = impl.hash_batch(m, samples, offsets, threads=3)
            streams = [impl.press(m, samples[int(offsets[r]):int(offsets[r + 1])])[1] for r in range(3)]
            assert sz.tolist() == [len(x) for x in streams]
            assert h.tolist() == [cpu_impl.fnv1a64(x) for x in streams]
            oo = np.concatenate([[0], np.cumsum([len(x) for x in streams])]).astype(np.uint64)
            cont = np.frombuffer(b"".join(streams), dtype=np.uint8)
            assert impl.hash_slices(cont, oo, threads=2).tolist() == h.tolist()


def test_building_blocks(oracle):
    import ctypes as C
    lib = oracle.lib
    lib.oracle_zigzag16.restype = C.c_uint16
    lib.oracle_zigzag16.argtypes = [C.c_int16]
    lib.oracle_unzigzag16.restype = C.c_int16
    lib.oracle_unzigzag16.argtypes = [C.c_uint16]
    assert [lib.oracle_zigzag16(x) for x in (0, -1, 1, -2, 2, 32767, -32768)] == \
        [0, 1, 2, 3, 4, 65534, 65535]
    for z in (0, 1, 2, 255, 256, 65534, 65535):
        assert lib.oracle_zigzag16(lib.oracle_unzigzag16(z)) == z
    lib.oracle_find_qts.restype = C.c_uint8
    lib.oracle_find_qts.argtypes = [C.POINTER(C.c_int16), C.c_uint64, C.c_uint8]
    for vals, q in (([4, 8, 0, -12], 2), ([0, 0], 5), ([64, 128], 5), ([3], 0), ([-32768], 5)):
        a = np.array(vals, dtype=np.int16)
        assert lib.oracle_find_qts(a.ctypes.data_as(C.POINTER(C.c_int16)), a.size, 5) == q


def test_ex_zd_scratch_cap_failure(oracle, ref):
    """ex_zd.c:411 -- stream longer than 2n + 1024 makes the reference return -1."""
    rng = np.random.default_rng(7)
    s = rng.integers(-32768, 32768, 20000).astype(np.int16) | 1   # q = 0, ~all exceptions
    assert ref.press("ex_zd", s)[0] != 0
    assert oracle.press("ex_zd", s)[0] != 0


def test_cpu_driver_round_trips(oracle, three_reads):
    samples, offsets, _ = three_reads
    for m in ("svb16_zd", "ex_zd", "shuff_vbe21_zd"):
        res = oracle.run_batch(m, samples, offsets, threads=2)
        assert res["rc"] == 0 and res["mismatches"] == 0 and res["errors"] == 0
