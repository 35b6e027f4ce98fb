import os
import sys

import numpy as _np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """Our CPU restatement (oracle/liboracle.so), built on demand."""
    from oracle import cpu_impl
    cpu_impl.build(ref=True)
    o = cpu_impl.load("oracle")
    o.set_huffman_table(open(os.path.join(ROOT, "honours_b200", "data", "NA12878_zd.huffman"), "rb").read())
    return o


@pytest.fixture(scope="session")
def ref():
    """The reference's own code compiled in place (oracle/_ref); skipped where absent."""
    from oracle import cpu_impl
    cpu_impl.build(ref=True)
    if not cpu_impl.available("ref"):
        pytest.skip("oracle/_ref/libpress_ref.so not built (no /root/reference here)")
    r = cpu_impl.load("ref")
    r.set_huffman_table(open(os.path.join(ROOT, "honours_b200", "data", "NA12878_zd.huffman"), "rb").read())
    return r


@pytest.fixture(scope="session")
def checker(oracle):
    """The strongest CPU checker present: the reference's own code compiled in place
    (oracle/_ref/libpress_ref.so -- it travels to the GPU box) when it exists, else the
    restatement (which tests/test_oracle.py pins to it byte for byte)."""
    from oracle import cpu_impl
    if cpu_impl.available("ref"):
        r = cpu_impl.load("ref")
        r.set_huffman_table(open(os.path.join(ROOT, "honours_b200", "data", "NA12878_zd.huffman"), "rb").read())
        return r
    return oracle


def svb0124_stream(rng, n, p4):
    """a hand-made svb0124 stream (streamvbyte_0124_encode.c:8-59 layout) whose values take 0, 1, 2 or 4
    bytes: a 16-bit signal never makes the reference write a 4-byte value, but its decoder reads them"""
    codes = rng.choice(4, size=n, p=[0.1, 0.6 - p4, 0.3, p4]).astype(_np.uint8)
    keys = _np.zeros((n + 3) // 4, dtype=_np.uint8)
    for sh in range(4):
        part = codes[sh::4]
        keys[: part.size] |= part << (2 * sh)
    data = bytearray()
    for c in codes:
        nb = (0, 1, 2, 4)[c]
        b = bytearray(rng.integers(0, 256, nb, dtype=_np.uint8).tobytes())
        if nb:
            b[-1] |= 1          # (a nonzero top byte, as an encoder would have it)
        data += b
    return keys.tobytes() + bytes(data)


def narrow_reads(rng):
    """signals whose zd values are tiny or sparse: widths 0 .. 3 for uint-zd (threads that bring fewer
    than 32 bits: the packer's OR path), mostly-zero values for svb0124-zd (threads with under 4 bytes)"""
    out = []
    for n in (1, 2, 17, 4096, 4097, 9001, 20000):
        out.append(_np.full(n, 321, dtype=_np.int16))                                    # constant: zd = 0, width 0
        out.append((700 - _np.cumsum(rng.integers(0, 2, n))).astype(_np.int16))          # steps of 0 / -1: zd in {0, 1}
        out.append((100 + _np.cumsum(rng.integers(0, 2, n))).astype(_np.int16))          # steps of 0 / +1: zd in {0, 2}
        d = _np.zeros(n, dtype=_np.int64)
        k = rng.random(n) < 0.02
        d[k] = rng.integers(-300, 300, int(k.sum()))                                   # flat with rare jumps
        out.append((1000 + _np.cumsum(d)).astype(_np.int16))
        out.append((_np.arange(n) % 7 - 3).astype(_np.int16))                            # small sawtooth: width 3..4
    return out
