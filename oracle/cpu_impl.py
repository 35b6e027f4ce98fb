"""oracle/cpu_impl.py -- TEST INFRASTRUCTURE ONLY.

ctypes front end for the two CPU implementations of the hot path:

* ``load("oracle")`` -> oracle/liboracle.so  (our C restatement, kind "port")
* ``load("ref")``    -> oracle/_ref/libpress_ref.so (the reference's own code
  compiled in place by ``make -C oracle ref``, kind "reference")

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this.  The product never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_DIR = "/root/reference/press"

METHODS = {
    "svb16_zd": 0,        # press.h svb12_zd_*
    "svb_zd": 1,          # press.h svb_zd_*_16
    "vbz_svb16_zd": 2,    # pre-zstd half of zstd_svb12_zd_*
    "vbz_svb_zd": 3,      # pre-zstd half of zstd_svb_zd_*_16
    "vbe21_zd": 4,        # press.h vbe21_zd_*_16
    "vbsse21_zd": 5,      # press.h vbsse21_zd_*_16
    "ex_zd": 6,           # press.h hasgam_vbsse21_zdq_*_16
    "shuff_vbe21_zd": 7,  # press.h shuffman_vbe21_zd_*_16
    "shuff_vbsse21_zd": 8,  # press.h shuffman_vbsse21_zd_*_16
    "vbbe21_zd": 9,  # press.h vbbe21_zd_*_16
    "shuff_vbbe21_zd": 10,  # press.h shuffman_vbbe21_zd_*_16
    "svb0124_zd": 11,     # press.h svb0124_zd_*_16
    "uint_zd": 12,        # press.h uint_zd_*_16
}

_u8p = C.POINTER(C.c_uint8)
_i16p = C.POINTER(C.c_int16)
_u64p = C.POINTER(C.c_uint64)


def build(ref: bool = True) -> None:
    """Compile liboracle.so and, when the reference tree is present, _ref."""
    subprocess.run(["make", "-s", "-C", HERE], check=True)
    if ref and os.path.isdir(REFERENCE_DIR):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


def lib_path(kind: str) -> str:
    if kind == "oracle":
        return os.path.join(HERE, "liboracle.so")
    if kind == "ref":
        return os.path.join(HERE, "_ref", "libpress_ref.so")
    raise ValueError(kind)


def available(kind: str) -> bool:
    return os.path.exists(lib_path(kind))


class CpuImpl:
    """One CPU implementation behind the uniform bound/press/depress adapter."""

    def __init__(self, kind: str):
        self.kind = kind
        self.lib = C.CDLL(lib_path(kind))
        p = kind
        self._bound = getattr(self.lib, f"{p}_bound")
        self._bound.restype = C.c_uint64
        self._bound.argtypes = [C.c_int, C.c_uint64]
        self._press = getattr(self.lib, f"{p}_press")
        self._press.restype = C.c_int
        self._press.argtypes = [C.c_int, _i16p, C.c_uint64, _u8p, _u64p]
        self._depress = getattr(self.lib, f"{p}_depress")
        self._depress.restype = C.c_int
        self._depress.argtypes = [C.c_int, _u8p, C.c_uint64, C.c_uint64, _i16p, _u64p]
        self._set_table = getattr(self.lib, f"{p}_set_huffman_table")
        self._set_table.restype = C.c_int
        self._set_table.argtypes = [C.c_char_p, C.c_size_t]
        self._run = getattr(self.lib, f"{p}_run_batch")
        self._run.restype = C.c_int
        self._run.argtypes = [C.c_int, _i16p, _u64p, C.c_uint64, C.c_int,
                              C.POINTER(C.c_double)]
        self._table_keepalive = None

    def set_huffman_table(self, blob: bytes) -> None:
        self._table_keepalive = bytes(blob)
        if self._set_table(self._table_keepalive, len(blob)) != 0:
            raise RuntimeError("bad huffman table")

    def bound(self, method: str, n: int) -> int:
        return int(self._bound(METHODS[method], n))

    def press(self, method: str, sig: np.ndarray) -> tuple[int, bytes]:
        """-> (return code, stream bytes)"""
        sig = np.ascontiguousarray(sig, dtype=np.int16)
        n = sig.size
        cap = self.bound(method, n) + 8 * n + 64
        out = np.zeros(cap, dtype=np.uint8)
        nout = C.c_uint64(cap)
        rc = self._press(METHODS[method], sig.ctypes.data_as(_i16p), n,
                         out.ctypes.data_as(_u8p), C.byref(nout))
        return rc, (out[: nout.value].tobytes() if rc == 0 else b"")

    def depress(self, method: str, stream: bytes, n: int) -> tuple[int, np.ndarray]:
        """-> (return code, samples); ``n`` = expected sample count"""
        buf = np.zeros(len(stream) + 64, dtype=np.uint8)   # decoders over-read <= 16 B
        buf[: len(stream)] = np.frombuffer(stream, dtype=np.uint8)
        out = np.zeros(max(n, len(stream)) + 64, dtype=np.int16)
        nout = C.c_uint64(n)
        rc = self._depress(METHODS[method], buf.ctypes.data_as(_u8p), len(stream), n,
                           out.ctypes.data_as(_i16p), C.byref(nout))
        return rc, out[: nout.value].copy()

    def nosimd_press(self, sig: np.ndarray) -> bytes:
        """svb12_zd_nosimd_press (press.c:1655-1662), reference build only."""
        sig = np.ascontiguousarray(sig, dtype=np.int16)
        self.lib.ref_nosimd_bound.restype = C.c_uint64
        self.lib.ref_nosimd_bound.argtypes = [C.c_uint64]
        cap = int(self.lib.ref_nosimd_bound(sig.size)) + 64
        out = np.zeros(cap, dtype=np.uint8)
        nout = C.c_uint64(cap)
        self.lib.ref_nosimd_press.argtypes = [_i16p, C.c_uint64, _u8p, _u64p]
        self.lib.ref_nosimd_press(sig.ctypes.data_as(_i16p), sig.size, out.ctypes.data_as(_u8p), C.byref(nout))
        return out[: nout.value].tobytes()

    def nosimd_depress(self, stream: bytes, n: int) -> np.ndarray:
        """svb12_zd_nosimd_depress (press.c:1664-1674), reference build only."""
        buf = np.zeros(len(stream) + 64, dtype=np.uint8)
        buf[: len(stream)] = np.frombuffer(stream, dtype=np.uint8)
        out = np.zeros(n + 64, dtype=np.int16)
        nout = C.c_uint64(n)
        self.lib.ref_nosimd_depress.argtypes = [_u8p, C.c_uint64, _i16p, _u64p]
        self.lib.ref_nosimd_depress(buf.ctypes.data_as(_u8p), n, out.ctypes.data_as(_i16p), C.byref(nout))
        return out[: nout.value].copy()

    def run_batch(self, method: str, samples: np.ndarray, offsets: np.ndarray,
                  threads: int = 1) -> dict:
        """press/test.c-style loop over the reads; wall-clock seconds per phase."""
        samples = np.ascontiguousarray(samples, dtype=np.int16)
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        res = (C.c_double * 5)()
        rc = self._run(METHODS[method], samples.ctypes.data_as(_i16p),
                       offsets.ctypes.data_as(_u64p), offsets.size - 1, threads, res)
        return {"rc": rc, "press_s": res[0], "depress_s": res[1],
                "bytes": int(res[2]), "mismatches": int(res[3]), "errors": int(res[4])}


def _hash_batch(self, method: str, samples: np.ndarray, offsets: np.ndarray, threads: int = 0):
    """-> (hashes uint64[nreads], sizes uint64[nreads]): FNV-1a-64 and length of every read's
    stream as this implementation writes it (size 0: press failed or empty read)."""
    samples = np.ascontiguousarray(samples, dtype=np.int16)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    nreads = offsets.size - 1
    hashes = np.zeros(nreads, dtype=np.uint64)
    sizes = np.zeros(nreads, dtype=np.uint64)
    f = getattr(self.lib, f"{self.kind}_hash_batch")
    f.restype = C.c_int
    f.argtypes = [C.c_int, _i16p, _u64p, C.c_uint64, C.c_int, _u64p, _u64p]
    f(METHODS[method], samples.ctypes.data_as(_i16p), offsets.ctypes.data_as(_u64p), nreads,
      threads or (os.cpu_count() or 1), hashes.ctypes.data_as(_u64p), sizes.ctypes.data_as(_u64p))
    return hashes, sizes


def _hash_slices(self, container: np.ndarray, oo: np.ndarray, threads: int = 0) -> np.ndarray:
    """FNV-1a-64 of container[oo[r]:oo[r + 1]] for every r."""
    container = np.ascontiguousarray(container, dtype=np.uint8)
    oo = np.ascontiguousarray(oo, dtype=np.uint64)
    nreads = oo.size - 1
    hashes = np.zeros(nreads, dtype=np.uint64)
    f = getattr(self.lib, f"{self.kind}_hash_slices")
    f.restype = C.c_int
    f.argtypes = [_u8p, _u64p, C.c_uint64, C.c_int, _u64p]
    f(container.ctypes.data_as(_u8p), oo.ctypes.data_as(_u64p), nreads, threads or (os.cpu_count() or 1),
      hashes.ctypes.data_as(_u64p))
    return hashes


CpuImpl.hash_batch = _hash_batch
CpuImpl.hash_slices = _hash_slices

_cache: dict[str, CpuImpl] = {}


def load(kind: str) -> CpuImpl:
    if kind not in _cache:
        _cache[kind] = CpuImpl(kind)
    return _cache[kind]


def fnv1a64(data: bytes) -> int:
    """FNV-1a 64-bit, the stream fingerprint used in SURVEY.md 8(c)."""
    h = 0xCBF29CE484222325
    arr = np.frombuffer(data, dtype=np.uint8)
    for b in arr.tolist():
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return h
