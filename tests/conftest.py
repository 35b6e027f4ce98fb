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
