import os
import sys

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
