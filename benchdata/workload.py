"""Synthetic NA12878-like read batches: lognormal read lengths
matching the NA12878 median / mean (thesis/plots/n-tab.tex), first samples
~ N(475, 35) (rawsig-tab.tex), deltas drawn from the one-byte zd histogram of
press/gen_huffman.c:11 with exception rate ``p_ex``.

Lengths / first samples / shifts are small per-read tables made here with
numpy; the samples themselves come from the integer-only generator of
benchdata/synth_core.h, on the device (libsynth_gpu.so) for the bench and on the
host (libsynth_host.so, no CUDA) for the CPU baseline and the tests -- both give
identical bits for the same (seed, read index).  Nothing here touches the
product library.
"""
from __future__ import annotations

import ctypes as C
import json
import math
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))

SEED = 0x5EED0001
MEDIAN_LEN = 80304.5
MEAN_LEN = 113471.4
MIN_LEN = 2024
MAX_LEN = 5_724_000
P_EX = 4.8546 / 113471.4          # exceptions per sample on NA12878 (ex-tab.tex)

MU = math.log(MEDIAN_LEN)
SIGMA = math.sqrt(2.0 * (math.log(MEAN_LEN) - MU))


def histogram() -> np.ndarray:
    with open(os.path.join(PKG, "NA12878_zd_hist.json")) as f:
        return np.asarray(json.load(f)["counts"], dtype=np.uint64)


def read_table(nreads: int, seed: int = SEED, first_read: int = 0, q_fraction: float = 0.0):
    """Per-read tables for reads [first_read, first_read + nreads) of the batch:
    (lengths int64, x0 int16, shift uint8).  Read r's row depends only on
    (seed, r), so any shard of the batch can be produced independently."""
    idx = np.arange(first_read, first_read + nreads, dtype=np.uint64)
    # counter-based: one Philox stream per block of 4096 reads keeps this cheap and shardable
    lengths = np.empty(nreads, dtype=np.int64)
    x0 = np.empty(nreads, dtype=np.int16)
    shift = np.zeros(nreads, dtype=np.uint8)
    blk = 4096
    b0, b1 = first_read // blk, (first_read + nreads + blk - 1) // blk
    for b in range(b0, b1):
        rng = np.random.Generator(np.random.Philox(key=seed, counter=[b, 0, 0, 0]))
        z = rng.standard_normal(blk)
        zx = rng.standard_normal(blk)
        u = rng.random(blk)
        lo, hi = max(first_read, b * blk), min(first_read + nreads, (b + 1) * blk)
        sl = slice(lo - b * blk, hi - b * blk)
        dst = slice(lo - first_read, hi - first_read)
        lengths[dst] = np.clip(np.rint(np.exp(MU + SIGMA * z[sl])), MIN_LEN, MAX_LEN).astype(np.int64)
        x0[dst] = np.clip(np.rint(475.0 + 35.0 * zx[sl]), 0, 2047).astype(np.int16)
        if q_fraction > 0:
            shift[dst] = np.where(u[sl] < q_fraction, 2, 0).astype(np.uint8)
    del idx
    return lengths, x0, shift


def offsets_of(lengths: np.ndarray) -> np.ndarray:
    off = np.zeros(lengths.size + 1, dtype=np.uint64)
    np.cumsum(lengths, out=off[1:])
    return off


def p_ex_u32(p_ex: float) -> int:
    return min(0xFFFFFFFF, int(round(p_ex * 2.0 ** 32)))


_libs: dict[str, C.CDLL] = {}


def _lib(which: str) -> C.CDLL:
    if which not in _libs:
        p = os.path.join(PKG, f"libsynth_{which}.so")
        if not os.path.exists(p):
            from . import build
            build.build()
        L = C.CDLL(p)
        vp, u64, u32 = C.c_void_p, C.c_uint64, C.c_uint32
        if which == "gpu":
            L.synth_gpu_device.restype = C.c_int
            L.synth_gpu_device.argtypes = [vp, vp, u64, vp, vp, vp, u32, u64, u64, vp]
            L.bench_link_probe.restype = C.c_int
            L.bench_link_probe.argtypes = [C.c_size_t, C.c_size_t, C.c_int, C.POINTER(C.c_double)]
        else:
            L.synth_host_fill.restype = C.c_int
            L.synth_host_fill.argtypes = [vp, vp, u64, vp, vp, vp, u32, u64, u64]
        _libs[which] = L
    return _libs[which]


def synth_host(offsets: np.ndarray, x0: np.ndarray, shift: np.ndarray | None = None,
               p_ex: float = P_EX, seed: int = SEED, first_read: int = 0) -> np.ndarray:
    """Samples of the reads described by ``offsets`` (uint64[nreads+1]) on the host.
    ``first_read`` is the batch index of offsets' read 0: read r's samples depend
    only on (seed, first_read + r)."""
    L = _lib("host")
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    x0 = np.ascontiguousarray(x0, dtype=np.int16)
    sh = None if shift is None else np.ascontiguousarray(shift, dtype=np.uint8)
    hist = histogram()
    out = np.empty(int(offsets[-1]), dtype=np.int16)
    rc = L.synth_host_fill(out.ctypes.data, offsets.ctypes.data, offsets.size - 1,
                                x0.ctypes.data, None if sh is None else sh.ctypes.data,
                                hist.ctypes.data, p_ex_u32(p_ex), seed, first_read)
    if rc != 0:
        raise RuntimeError("synth_host_fill failed")
    return out


def synth_device(d_samples: int, d_offsets: int, nreads: int, d_x0: int, d_shift: int | None,
                 p_ex: float = P_EX, seed: int = SEED, first_read: int = 0, stream: int = 0) -> None:
    """Fill device memory (raw pointers) with the same samples synth_host returns."""
    L = _lib("gpu")
    hist = histogram()
    rc = L.synth_gpu_device(d_samples, d_offsets, nreads, d_x0, d_shift, hist.ctypes.data,
                                  p_ex_u32(p_ex), seed, first_read, stream)
    if rc != 0:
        raise RuntimeError("synth_gpu_device failed (no CUDA device?)")


def link_probe(total_bytes: int = 1 << 30, chunk_bytes: int = 64 << 20, reps: int = 3) -> dict:
    """Plain pinned cudaMemcpyAsync bandwidth of this process's GPU (GB/s): each direction alone
    and both at once -- the ceiling the host-buffer (e2e) path is measured against."""
    out = (C.c_double * 4)()
    if _lib("gpu").bench_link_probe(total_bytes, chunk_bytes, reps, out) != 0:
        raise RuntimeError("bench_link_probe failed")
    return {"h2d_gbs": out[0], "d2h_gbs": out[1], "h2d_duplex_gbs": out[2], "d2h_duplex_gbs": out[3],
            "bytes": total_bytes, "chunk_bytes": chunk_bytes, "reps": reps}
