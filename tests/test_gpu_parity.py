"""GPU parity tests (run with -m gpu on a B200): every stream the CUDA path
writes must be byte-identical to the oracle's (== the reference's) stream for
the same read, and every decode must return the original samples.  All calls
go through the C ABI (include/press_gpu.h) via honours_b200.press."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden")
IMPLEMENTED = ["svb16_zd", "svb_zd", "vbz_svb16_zd", "vbz_svb_zd"]


@pytest.fixture(scope="module")
def gpu():
    from honours_b200 import press
    press.init(0)
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
def test_three_reads_per_read_dropin(gpu, oracle, three_reads, method):
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
        assert st == oracle.press(method, s)[1]
        rc, back = gpu.depress(method, st, s.size)
        assert rc == 0 and np.array_equal(back, s)


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_edge_lengths_per_read(gpu, oracle, method):
    rng = np.random.default_rng(42)
    for n in [1, 2, 3, 7, 8, 9, 15, 16, 17, 63, 64, 65, 4095, 4096, 4097, 8191, 8192, 8193, 12289]:
        for kind in ("walk", "noise"):
            s = rand_read(rng, kind, n)
            rc, st = gpu.press(method, s)
            ro, so = oracle.press(method, s)
            assert rc == 0 and ro == 0 and st == so, (method, kind, n)
            rc, back = gpu.depress(method, so, n)
            assert rc == 0 and np.array_equal(back, s), (method, kind, n)


def make_batch(rng, nreads, lo, hi, kinds=("walk", "spiky", "q", "noise")):
    lens = rng.integers(lo, hi, nreads)
    reads = [rand_read(rng, kinds[i % len(kinds)], int(n)) for i, n in enumerate(lens)]
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    return np.concatenate(reads), offsets, reads


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_batch_host_matches_oracle(gpu, oracle, method):
    """tight container: stream r at out_offsets[r], all byte-identical to the oracle;
    reads are back to back so most start at a misaligned address."""
    rng = np.random.default_rng(7)
    samples, offsets, reads = make_batch(rng, 300, 1, 30000)
    rc, cont, oo = gpu.batch_press_host(method, samples, offsets)
    assert rc == 0
    expect = [oracle.press(method, s)[1] for s in reads]
    assert int(oo[0]) == 0
    assert np.diff(oo.astype(np.int64)).tolist() == [len(e) for e in expect]
    assert cont.tobytes() == b"".join(expect)
    back = gpu.batch_depress_host(method, cont, oo, offsets)
    assert np.array_equal(back, samples)


@pytest.mark.parametrize("method", IMPLEMENTED)
def test_batch_with_empty_and_tiny_reads(gpu, oracle, method):
    rng = np.random.default_rng(9)
    lens = [0, 1, 0, 5, 4096, 0, 4097, 8, 0]
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
    for method in ("svb16_zd", "vbz_svb_zd"):
        rc, st = gpu.press(method, s)
        assert rc == 0 and st == oracle.press(method, s)[1]
        rc, back = gpu.depress(method, st, s.size)
        assert rc == 0 and np.array_equal(back, s)


def test_device_batch_api_torch_pointers(gpu, oracle):
    """the measured path: device-resident in/out, raw pointers through the C ABI"""
    import torch
    rng = np.random.default_rng(5)
    samples, offsets, reads = make_batch(rng, 2000, 2000, 60000, kinds=("walk", "spiky"))
    method = "svb16_zd"
    dev = torch.device("cuda:0")
    d_s = torch.from_numpy(samples).to(dev)
    d_off = torch.from_numpy(offsets.astype(np.int64)).to(dev)
    nreads, ns = len(reads), samples.size
    cap = gpu.batch_bound(method, nreads, ns)
    d_out = torch.empty(cap, dtype=torch.uint8, device=dev)
    d_oo = torch.empty(nreads + 1, dtype=torch.int64, device=dev)
    sb = gpu.batch_scratch_bytes(method, nreads, ns)
    d_scr = torch.empty(sb, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    gpu.batch_press(method, d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_out.data_ptr(), cap,
                    d_oo.data_ptr(), d_scr.data_ptr(), sb, stream)
    assert gpu.batch_status(d_scr.data_ptr(), stream) == 0
    oo = d_oo.cpu().numpy()
    total = int(oo[-1])
    expect = b"".join(oracle.press(method, s)[1] for s in reads)
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
