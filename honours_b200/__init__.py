"""honours_b200 -- B200 (sm_100a) implementation of the `press` zd / svb16 / svb32 /
exception / static-Huffman hot path (sashajenner/honours, press/), behind the
reference's own press.h method interface.

The product is honours_b200/libpress_gpu.so (hand-written CUDA + a C ABI,
include/press_gpu.h).  `honours_b200.press` is the thin Python mirror of that
ABI used by the tests and bench.py; it never computes anything itself and it
raises if the CUDA library is missing -- there is no CPU fallback.
"""
from .press import (  # noqa: F401
    METHODS,
    PressGpuError,
    lib,
    lib_path,
)

__version__ = "0.1"
