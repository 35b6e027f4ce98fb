"""press_gpu_svb16_zd_press_static: the entry point that allocates nothing (its scratch is a static
device array inside the library, no counterpart in
the reference).  Its container must be byte-identical to press_gpu_batch_press's on the same batch,
on two calls in a row (the array keeps the first call's contents), with no device allocation
between the calls, and batches beyond its limits must be rejected.  Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    from honours_b200 import press
    press.init(0)
    return press


def _batch(seed, lens):
    rng = np.random.default_rng(seed)
    reads = []
    for i, n in enumerate(lens):
        step = 6 if i % 3 else 300          # every third read has two-byte deltas
        reads.append((400 + np.cumsum(rng.integers(-step, step + 1, n))).astype(np.int16))
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    return (np.concatenate(reads) if reads else np.zeros(0, np.int16)), off


def test_static_scratch_press_matches_batch_press(gpu):
    import torch
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    batches = [_batch(1, [70001, 1, 0, 9, 8192, 8193, 123457, 40000, 2, 300000]),
               _batch(2, [5, 250000, 8191, 33, 0, 99999])]       # second call: fewer tiles, other data
    # everything either call needs is allocated up front
    bufs = []
    for samples, off in batches:
        nreads, ns = off.size - 1, int(off[-1])
        cap = gpu.batch_bound("svb16_zd", nreads, ns)
        sb = gpu.batch_scratch_bytes("svb16_zd", nreads, ns)
        bufs.append(dict(
            nreads=nreads, ns=ns, cap=cap, sb=sb,
            d_s=torch.from_numpy(samples).to(dev), d_off=torch.from_numpy(off.astype(np.int64)).to(dev),
            d_out=torch.zeros(cap + 64, dtype=torch.uint8, device=dev), d_oo=torch.empty(nreads + 1, dtype=torch.int64, device=dev),
            d_scr=torch.empty(sb, dtype=torch.uint8, device=dev),
            s_out=torch.zeros(cap + 64, dtype=torch.uint8, device=dev), s_oo=torch.empty(nreads + 1, dtype=torch.int64, device=dev),
            s_status=torch.full((1,), 77, dtype=torch.int32, device=dev)))
    # the existing entry point
    for b in bufs:
        gpu.batch_press("svb16_zd", b["d_s"].data_ptr(), b["d_off"].data_ptr(), b["nreads"], b["ns"], b["d_out"].data_ptr(),
                        b["cap"], b["d_oo"].data_ptr(), b["d_scr"].data_ptr(), b["sb"], st)
        assert gpu.batch_status(b["d_scr"].data_ptr(), st) == 0
    # one warm-up call (one-time set-up), then two calls in a row with no allocation in between
    w = bufs[1]
    gpu.svb16_zd_press_static(w["d_s"].data_ptr(), w["d_off"].data_ptr(), w["nreads"], w["ns"], w["s_out"].data_ptr(),
                              w["cap"], w["s_oo"].data_ptr(), w["s_status"].data_ptr(), st)
    torch.cuda.synchronize()
    w["s_out"].zero_()
    free0 = torch.cuda.mem_get_info()[0]
    for b in bufs:
        gpu.svb16_zd_press_static(b["d_s"].data_ptr(), b["d_off"].data_ptr(), b["nreads"], b["ns"], b["s_out"].data_ptr(),
                                  b["cap"], b["s_oo"].data_ptr(), b["s_status"].data_ptr(), st)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "the static-scratch call changed the device's free memory"
    for b in bufs:
        assert int(b["s_status"].item()) == 0
        assert torch.equal(b["s_oo"], b["d_oo"])
        total = int(b["d_oo"][-1].item())
        assert total > 0 and torch.equal(b["s_out"][:total], b["d_out"][:total])       # bit-exact (byte work)
        assert bool((b["s_out"][total:] == 0).all())


def test_static_scratch_reports_capacity_and_rejects_large_batches(gpu):
    import torch
    from honours_b200 import press
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    samples, off = _batch(3, [50000, 60000])
    d_s = torch.from_numpy(samples).to(dev)
    d_off = torch.from_numpy(off.astype(np.int64)).to(dev)
    cap = gpu.batch_bound("svb16_zd", 2, int(off[-1]))
    d_out = torch.zeros(cap + 64, dtype=torch.uint8, device=dev)      # (full size: only out_capacity is short)
    d_oo = torch.empty(3, dtype=torch.int64, device=dev)
    d_status = torch.zeros(1, dtype=torch.int32, device=dev)
    gpu.svb16_zd_press_static(d_s.data_ptr(), d_off.data_ptr(), 2, int(off[-1]), d_out.data_ptr(), 1024, d_oo.data_ptr(),
                              d_status.data_ptr(), st)
    torch.cuda.synchronize()
    assert int(d_status.item()) & 1                                 # capacity exceeded, as press_gpu_batch_status reports it
    # ... and the next call starts from a clean error word
    d_big = torch.zeros(cap + 64, dtype=torch.uint8, device=dev)
    gpu.svb16_zd_press_static(d_s.data_ptr(), d_off.data_ptr(), 2, int(off[-1]), d_big.data_ptr(), cap, d_oo.data_ptr(),
                              d_status.data_ptr(), st)
    torch.cuda.synchronize()
    assert int(d_status.item()) == 0
    for nreads, ns in ((press.STATIC_MAX_READS + 1, 1000), (2, press.STATIC_MAX_SAMPLES + 1)):
        with pytest.raises(press.PressGpuError, match="exceed the static scratch's limits"):
            gpu.svb16_zd_press_static(d_s.data_ptr(), d_off.data_ptr(), nreads, ns, d_big.data_ptr(), cap, d_oo.data_ptr(), 0, st)
