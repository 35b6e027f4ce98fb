"""ctypes mirror of include/press_gpu.h.

Per-read calls mirror the reference's press.h triples
(``X_bound(nin)``, ``X_press(in, nin, out, *nout)``, ``X_depress(in, nin, out, *nout)``,
press/press.h:10-33) with numpy arrays as the host buffers; the batch calls take
raw device pointers (``tensor.data_ptr()``) exactly like the C ABI.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))

METHODS = {
    "svb16_zd": 0,
    "svb_zd": 1,
    "vbz_svb16_zd": 2,
    "vbz_svb_zd": 3,
    "vbe21_zd": 4,
    "vbsse21_zd": 5,
    "ex_zd": 6,
    "shuff_vbe21_zd": 7,
    "shuff_vbsse21_zd": 8,
    "vbbe21_zd": 9,
    "shuff_vbbe21_zd": 10,
    "svb0124_zd": 11,
    "uint_zd": 12,
}

# per-read entry points of include/press_gpu.h, by method
PER_READ = {
    "svb16_zd": ("svb12_zd_gpu_bound", "svb12_zd_gpu_press", "svb12_zd_gpu_depress"),
    "svb_zd": ("svb_zd_gpu_bound_16", "svb_zd_gpu_press_16", "svb_zd_gpu_depress_16"),
    "vbz_svb16_zd": ("vbz_svb12_zd_gpu_bound", "vbz_svb12_zd_gpu_press", "vbz_svb12_zd_gpu_depress"),
    "vbz_svb_zd": ("vbz_svb_zd_gpu_bound_16", "vbz_svb_zd_gpu_press_16", "vbz_svb_zd_gpu_depress_16"),
    "vbe21_zd": ("vbe21_zd_gpu_bound_16", "vbe21_zd_gpu_press_16", "vbe21_zd_gpu_depress_16"),
    "vbsse21_zd": ("vbsse21_zd_gpu_bound_16", "vbsse21_zd_gpu_press_16", "vbsse21_zd_gpu_depress_16"),
    "ex_zd": ("hasgam_vbsse21_zdq_gpu_bound_16", "hasgam_vbsse21_zdq_gpu_press_16",
              "hasgam_vbsse21_zdq_gpu_depress_16"),
    "shuff_vbe21_zd": ("shuffman_vbe21_zd_gpu_bound_16", "shuffman_vbe21_zd_gpu_press_16",
                       "shuffman_vbe21_zd_gpu_depress_16"),
    "shuff_vbsse21_zd": ("shuffman_vbsse21_zd_gpu_bound_16", "shuffman_vbsse21_zd_gpu_press_16",
                         "shuffman_vbsse21_zd_gpu_depress_16"),
    "vbbe21_zd": ("vbbe21_zd_gpu_bound_16", "vbbe21_zd_gpu_press_16", "vbbe21_zd_gpu_depress_16"),
    "shuff_vbbe21_zd": ("shuffman_vbbe21_zd_gpu_bound_16", "shuffman_vbbe21_zd_gpu_press_16",
                        "shuffman_vbbe21_zd_gpu_depress_16"),
    "svb0124_zd": ("svb0124_zd_gpu_bound_16", "svb0124_zd_gpu_press_16", "svb0124_zd_gpu_depress_16"),
    "uint_zd": ("uint_zd_gpu_bound_16", "uint_zd_gpu_press_16", "uint_zd_gpu_depress_16"),
}


class PressGpuError(RuntimeError):
    pass


def lib_path() -> str:
    # PRESS_GPU_LIB: a variant build of the same library (tuning experiments, tools/)
    return os.environ.get("PRESS_GPU_LIB") or os.path.join(PKG, "libpress_gpu.so")


_lib = None


def lib() -> C.CDLL:
    """Load libpress_gpu.so.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise PressGpuError(
                f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        L = C.CDLL(p)
        vp, u64, i32, sz = C.c_void_p, C.c_uint64, C.c_int, C.c_size_t
        L.press_gpu_init.restype = i32
        L.press_gpu_init.argtypes = [i32]
        L.press_gpu_last_error.restype = C.c_char_p
        L.press_gpu_version.restype = C.c_char_p
        L.press_gpu_launch_count.restype = u64
        L.press_gpu_method_implemented.restype = i32
        L.press_gpu_method_implemented.argtypes = [i32]
        L.press_gpu_set_huffman_table.restype = i32
        L.press_gpu_set_huffman_table.argtypes = [C.c_char_p, sz]
        L.press_gpu_batch_bound.restype = u64
        L.press_gpu_batch_bound.argtypes = [i32, u64, u64]
        L.press_gpu_batch_scratch_bytes.restype = sz
        L.press_gpu_batch_scratch_bytes.argtypes = [i32, u64, u64]
        L.press_gpu_set_exception_capacity.restype = None
        L.press_gpu_set_exception_capacity.argtypes = [C.c_double]
        L.press_gpu_batch_press.restype = i32
        L.press_gpu_batch_press.argtypes = [i32, vp, vp, u64, u64, vp, u64, vp, vp, sz, vp]
        L.press_gpu_batch_depress.restype = i32
        L.press_gpu_batch_depress.argtypes = [i32, vp, vp, u64, vp, u64, u64, vp, vp, sz, vp]
        L.press_gpu_sidecar_bytes.restype = sz
        L.press_gpu_sidecar_bytes.argtypes = [i32, u64, u64]
        L.press_gpu_batch_press_sidecar.restype = i32
        L.press_gpu_batch_press_sidecar.argtypes = [i32, vp, vp, u64, u64, vp, u64, vp, vp, sz, vp, sz, vp]
        L.press_gpu_batch_depress_sidecar.restype = i32
        L.press_gpu_batch_depress_sidecar.argtypes = [i32, vp, vp, u64, vp, u64, u64, vp, vp, sz, vp, sz, vp]
        L.press_gpu_batch_status.restype = i32
        L.press_gpu_batch_status.argtypes = [vp, vp]
        L.press_gpu_svb16_zd_press_static.restype = i32
        L.press_gpu_svb16_zd_press_static.argtypes = [vp, vp, u64, u64, vp, u64, vp, vp, vp]
        L.press_gpu_batch_press_host.restype = i32
        L.press_gpu_batch_press_host.argtypes = [i32, vp, vp, u64, vp, u64, vp]
        L.press_gpu_batch_depress_host.restype = i32
        L.press_gpu_batch_depress_host.argtypes = [i32, vp, vp, vp, u64, vp]
        L.press_gpu_global_offsets.restype = i32
        L.press_gpu_global_offsets.argtypes = [vp, i32, u64, vp, vp]
        L.press_gpu_comm_unique_id.restype = i32
        L.press_gpu_comm_unique_id.argtypes = [vp]
        L.press_gpu_comm_init.restype = i32
        L.press_gpu_comm_init.argtypes = [vp, i32, i32]
        L.press_gpu_comm_destroy.restype = None
        L.press_gpu_allgather_offsets.restype = i32
        L.press_gpu_allgather_offsets.argtypes = [vp, u64, vp, vp, vp]
        L.press_gpu_blow5_open.restype = vp
        L.press_gpu_blow5_open.argtypes = [C.c_char_p]
        L.press_gpu_blow5_close.argtypes = [vp]
        L.press_gpu_blow5_signal_press.restype = i32
        L.press_gpu_blow5_signal_press.argtypes = [vp]
        L.press_gpu_blow5_next_batch.restype = C.c_int64
        L.press_gpu_blow5_next_batch.argtypes = [vp, u64, u64, vp, vp, vp, sz]
        L.press_gpu_container_write.restype = i32
        L.press_gpu_container_write.argtypes = [C.c_char_p, i32, u64, vp, vp, vp]
        L.press_gpu_container_info.restype = i32
        L.press_gpu_container_info.argtypes = [C.c_char_p, C.POINTER(i32), C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]
        L.press_gpu_container_read.restype = i32
        L.press_gpu_container_read.argtypes = [C.c_char_p, vp, vp, vp]
        L.press_gpu_profile_enable.restype = None
        L.press_gpu_profile_enable.argtypes = [i32]
        L.press_gpu_profile_collect.restype = i32
        L.press_gpu_profile_collect.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        L.press_gpu_host_alloc.restype = vp
        L.press_gpu_host_alloc.argtypes = [sz]
        L.press_gpu_host_free.argtypes = [vp]
        _lib = L
    return _lib


def implemented_methods() -> list[str]:
    L = lib()
    return [m for m, k in METHODS.items() if L.press_gpu_method_implemented(k)]


def launch_count() -> int:
    return int(lib().press_gpu_launch_count())


def last_error() -> str:
    return lib().press_gpu_last_error().decode()


def init(device: int = -1) -> None:
    if lib().press_gpu_init(device) != 0:
        raise PressGpuError(last_error())


def set_huffman_table(blob: bytes | None = None) -> None:
    if blob is None:
        blob = open(os.path.join(PKG, "data", "NA12878_zd.huffman"), "rb").read()
    if lib().press_gpu_set_huffman_table(blob, len(blob)) != 0:
        raise PressGpuError(last_error())


def set_exception_capacity(entries_per_sample: float) -> None:
    """Exception pool size used by batch_scratch_bytes() for methods 4-7."""
    lib().press_gpu_set_exception_capacity(float(entries_per_sample))


# ---------------------------------------------------------------------------
# per-read mirror of press.h

def bound(method: str, nin: int) -> int:
    f = getattr(lib(), PER_READ[method][0])
    f.restype = C.c_uint64
    f.argtypes = [C.c_uint64 if method in ("svb16_zd", "svb_zd", "svb0124_zd", "uint_zd") else C.c_uint32]
    return int(f(nin))


def press(method: str, sig: np.ndarray) -> tuple[int, bytes]:
    """``X_press(in, nin, out, &nout)`` -> (return code, stream bytes[:nout])."""
    L = lib()
    sig = np.ascontiguousarray(sig, dtype=np.int16)
    n = sig.size
    cap = bound(method, n) + 8 * n + 64     # harness slack, like oracle/cpu_driver.c
    out = np.zeros(cap, dtype=np.uint8)
    nout = C.c_uint64(cap)
    f = getattr(L, PER_READ[method][1])
    rc = 0
    if method in ("svb16_zd", "svb_zd", "svb0124_zd"):
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64)]
        f(sig.ctypes.data, n, out.ctypes.data, C.byref(nout))
    elif method == "uint_zd":
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64)]
        rc = f(sig.ctypes.data, n, out.ctypes.data, C.byref(nout))
    elif method in ("vbe21_zd", "vbsse21_zd", "vbbe21_zd"):
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]
        f(sig.ctypes.data, n, out.ctypes.data, C.byref(nout))
    elif method in ("shuff_vbe21_zd", "shuff_vbsse21_zd", "shuff_vbbe21_zd"):
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]
        rc = f(None, sig.ctypes.data, n, out.ctypes.data, C.byref(nout))
    else:
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]
        rc = f(sig.ctypes.data, n, out.ctypes.data, C.byref(nout))
    return rc, out[: nout.value].tobytes()


def depress(method: str, stream: bytes, n: int) -> tuple[int, np.ndarray]:
    """``X_depress(in, nin, out, &nout)`` -> (return code, samples[:nout]).

    ``nin`` is the sample count for the svb methods and the byte count for the
    others, as in press.h."""
    L = lib()
    buf = np.zeros(len(stream) + 64, dtype=np.uint8)
    buf[: len(stream)] = np.frombuffer(stream, dtype=np.uint8)
    out = np.zeros(n + 64, dtype=np.int16)
    f = getattr(L, PER_READ[method][2])
    rc = 0
    if method in ("svb16_zd", "svb_zd", "svb0124_zd"):
        nout = C.c_uint64(n)
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64)]
        f(buf.ctypes.data, n, out.ctypes.data, C.byref(nout))
    elif method == "uint_zd":
        nout = C.c_uint64(n)
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64)]
        rc = f(buf.ctypes.data, n, out.ctypes.data, C.byref(nout))
    elif method in ("vbe21_zd", "vbsse21_zd", "vbbe21_zd"):
        nout = C.c_uint32(n)
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint32)]
        f(buf.ctypes.data, len(stream), out.ctypes.data, C.byref(nout))
    elif method in ("shuff_vbe21_zd", "shuff_vbsse21_zd", "shuff_vbbe21_zd"):
        nout = C.c_uint32(n)
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint32)]
        rc = f(None, buf.ctypes.data, len(stream), out.ctypes.data, C.byref(nout))
    else:
        nout = C.c_uint32(n)
        f.restype = C.c_int
        f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint32)]
        rc = f(buf.ctypes.data, len(stream), out.ctypes.data, C.byref(nout))
    return rc, out[: nout.value].copy()


# ---------------------------------------------------------------------------
# batch mirror (raw pointers; host variants take numpy arrays)

def batch_bound(method: str, nreads: int, nsamples: int) -> int:
    return int(lib().press_gpu_batch_bound(METHODS[method], nreads, nsamples))


def batch_scratch_bytes(method: str, nreads: int, nsamples: int) -> int:
    return int(lib().press_gpu_batch_scratch_bytes(METHODS[method], nreads, nsamples))


def sidecar_bytes(method: str, nreads: int, nsamples: int) -> int:
    """Bytes of the optional decode side-car of `method` for a batch (0: the method has none)."""
    return int(lib().press_gpu_sidecar_bytes(METHODS[method], nreads, nsamples))


def batch_press(method: str, d_samples: int, d_offsets: int, nreads: int, nsamples: int,
                d_out: int, out_capacity: int, d_out_offsets: int, d_scratch: int,
                scratch_bytes: int, stream: int = 0, sidecar: int = 0, sidecar_bytes: int = 0) -> None:
    if sidecar:
        rc = lib().press_gpu_batch_press_sidecar(METHODS[method], d_samples, d_offsets, nreads, nsamples,
                                                 d_out, out_capacity, d_out_offsets, d_scratch,
                                                 scratch_bytes, sidecar, sidecar_bytes, stream)
        if rc != 0:
            raise PressGpuError(last_error())
        return
    rc = lib().press_gpu_batch_press(METHODS[method], d_samples, d_offsets, nreads, nsamples,
                                     d_out, out_capacity, d_out_offsets, d_scratch,
                                     scratch_bytes, stream)
    if rc != 0:
        raise PressGpuError(last_error())


STATIC_MAX_READS = 4096             # include/press_gpu.h PRESS_GPU_STATIC_MAX_READS
STATIC_MAX_SAMPLES = 1 << 25        # PRESS_GPU_STATIC_MAX_SAMPLES


def svb16_zd_press_static(d_samples: int, d_offsets: int, nreads: int, nsamples: int, d_out: int,
                          out_capacity: int, d_out_offsets: int, d_status: int = 0, stream: int = 0) -> None:
    """press_gpu_svb16_zd_press_static: svb16-zd batch compression on the library's static device
    scratch (no allocation in the call; limits above; calls overlapping on the GPU race)."""
    rc = lib().press_gpu_svb16_zd_press_static(d_samples, d_offsets, nreads, nsamples, d_out, out_capacity,
                                               d_out_offsets, d_status, stream)
    if rc != 0:
        raise PressGpuError(last_error())


def batch_depress(method: str, d_in: int, d_in_offsets: int, in_bytes: int, d_offsets: int,
                  nreads: int, nsamples: int, d_samples_out: int, d_scratch: int,
                  scratch_bytes: int, stream: int = 0, sidecar: int = 0, sidecar_bytes: int = 0) -> None:
    if sidecar:
        rc = lib().press_gpu_batch_depress_sidecar(METHODS[method], d_in, d_in_offsets, in_bytes, d_offsets,
                                                   nreads, nsamples, d_samples_out, d_scratch,
                                                   scratch_bytes, sidecar, sidecar_bytes, stream)
        if rc != 0:
            raise PressGpuError(last_error())
        return
    rc = lib().press_gpu_batch_depress(METHODS[method], d_in, d_in_offsets, in_bytes, d_offsets,
                                       nreads, nsamples, d_samples_out, d_scratch,
                                       scratch_bytes, stream)
    if rc != 0:
        raise PressGpuError(last_error())


def batch_status(d_scratch: int, stream: int = 0) -> int:
    return int(lib().press_gpu_batch_status(d_scratch, stream))


def batch_press_host(method: str, samples: np.ndarray, offsets: np.ndarray):
    """Host-buffer batch compress -> (rc, container bytes (uint8 array), out_offsets)."""
    samples = np.ascontiguousarray(samples, dtype=np.int16)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    nreads = offsets.size - 1
    nsamples = int(offsets[-1] - offsets[0])
    cap = batch_bound(method, nreads, nsamples)
    out = np.empty(cap, dtype=np.uint8)
    oo = np.zeros(nreads + 1, dtype=np.uint64)
    rc = lib().press_gpu_batch_press_host(METHODS[method], samples.ctypes.data, offsets.ctypes.data,
                                          nreads, out.ctypes.data, cap, oo.ctypes.data)
    if rc < 0:
        raise PressGpuError(last_error())
    return rc, out[: int(oo[-1])], oo


def batch_depress_host(method: str, container: np.ndarray, in_offsets: np.ndarray,
                       offsets: np.ndarray) -> np.ndarray:
    container = np.ascontiguousarray(container, dtype=np.uint8)
    in_offsets = np.ascontiguousarray(in_offsets, dtype=np.uint64)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    nreads = offsets.size - 1
    out = np.zeros(int(offsets[-1]), dtype=np.int16)
    rc = lib().press_gpu_batch_depress_host(METHODS[method], container.ctypes.data,
                                            in_offsets.ctypes.data, offsets.ctypes.data,
                                            nreads, out.ctypes.data)
    if rc != 0:
        raise PressGpuError(last_error())
    return out


def global_offsets(d_gathered: int, nranks: int, nreads: int, d_global: int, stream: int = 0) -> None:
    if lib().press_gpu_global_offsets(d_gathered, nranks, nreads, d_global, stream) != 0:
        raise PressGpuError(last_error())


def comm_unique_id() -> bytes:
    buf = (C.c_uint8 * 128)()
    if lib().press_gpu_comm_unique_id(buf) != 0:
        raise PressGpuError(last_error())
    return bytes(buf)


def comm_init(uid: bytes, rank: int, nranks: int) -> None:
    buf = (C.c_uint8 * 128).from_buffer_copy(uid)
    if lib().press_gpu_comm_init(buf, rank, nranks) != 0:
        raise PressGpuError(last_error())


def comm_destroy() -> None:
    lib().press_gpu_comm_destroy()


def allgather_offsets(d_out_offsets: int, nreads: int, d_gathered: int, d_global: int, stream: int = 0) -> None:
    """The path's one exchange through the C ABI: ncclAllGather of the per-read byte positions +
    the container index (press_gpu_global_offsets) on the same stream."""
    if lib().press_gpu_allgather_offsets(d_out_offsets, nreads, d_gathered, d_global, stream) != 0:
        raise PressGpuError(last_error())


def profile_enable(on: bool) -> None:
    lib().press_gpu_profile_enable(1 if on else 0)


def profile_collect() -> tuple[list[float], list[int]]:
    """-> ([encode ms, decode ms], [encode launches, decode launches]) since the last call."""
    ms = (C.c_double * 2)()
    n = (C.c_uint64 * 2)()
    if lib().press_gpu_profile_collect(ms, n) != 0:
        raise PressGpuError(last_error())
    return [ms[0], ms[1]], [int(n[0]), int(n[1])]


def pinned_array(nbytes: int) -> np.ndarray:
    """uint8 numpy view of cudaHostAlloc'd memory (kept alive by the returned array's base)."""
    p = lib().press_gpu_host_alloc(nbytes)
    if not p:
        raise PressGpuError(last_error())
    buf = (C.c_uint8 * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=np.uint8)
    _pinned[arr.ctypes.data] = p
    return arr


def pinned_free(arr: np.ndarray) -> None:
    p = _pinned.pop(arr.ctypes.data, None)
    if p:
        lib().press_gpu_host_free(p)


_pinned: dict[int, int] = {}


# ---------------------------------------------------------------------------
# records in, container out

def blow5_read_all(path: str, max_reads: int = 1 << 16, max_samples: int = 1 << 26):
    """All records of a BLOW5 file through press_gpu_blow5_* -> (samples int16, offsets uint64, ids,
    signal_press).  svb-zd coded signals are decoded by the GPU inside the call."""
    L = lib()
    f = L.press_gpu_blow5_open(path.encode())
    if not f:
        raise PressGpuError(f"cannot open {path}")
    try:
        sp = L.press_gpu_blow5_signal_press(f)
        parts, lens, ids = [], [], []
        while True:
            s = np.zeros(max_samples, dtype=np.int16)
            o = np.zeros(max_reads + 1, dtype=np.uint64)
            idb = np.zeros(max_reads * 64, dtype=np.uint8)
            k = L.press_gpu_blow5_next_batch(f, max_reads, max_samples, s.ctypes.data, o.ctypes.data, idb.ctypes.data, 64)
            if k < 0:
                raise PressGpuError(last_error() or "blow5 read failed")
            if k == 0:
                break
            parts.append(s[: int(o[k])].copy())
            lens += np.diff(o[: k + 1].astype(np.int64)).tolist()
            ids += [bytes(idb[64 * i: 64 * i + 64]).split(b"\0")[0].decode() for i in range(k)]
        samples = np.concatenate(parts) if parts else np.zeros(0, np.int16)
        offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        return samples, offsets, ids, sp
    finally:
        L.press_gpu_blow5_close(f)


def container_write(path: str, method: str, offsets: np.ndarray, out_offsets: np.ndarray, container: np.ndarray) -> None:
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    out_offsets = np.ascontiguousarray(out_offsets, dtype=np.uint64)
    container = np.ascontiguousarray(container, dtype=np.uint8)
    if lib().press_gpu_container_write(path.encode(), METHODS[method], offsets.size - 1, offsets.ctypes.data,
                                       out_offsets.ctypes.data, container.ctypes.data) != 0:
        raise PressGpuError(f"cannot write {path}")


def container_read(path: str):
    """-> (method name, offsets, out_offsets, container bytes)"""
    L = lib()
    m, nr, ns, nb = C.c_int(), C.c_uint64(), C.c_uint64(), C.c_uint64()
    if L.press_gpu_container_info(path.encode(), C.byref(m), C.byref(nr), C.byref(ns), C.byref(nb)) != 0:
        raise PressGpuError(f"{path}: not a container")
    off = np.zeros(nr.value + 1, dtype=np.uint64)
    oo = np.zeros(nr.value + 1, dtype=np.uint64)
    cont = np.zeros(nb.value + 64, dtype=np.uint8)
    if L.press_gpu_container_read(path.encode(), off.ctypes.data, oo.ctypes.data, cont.ctypes.data) != 0:
        raise PressGpuError(f"{path}: truncated container")
    name = {v: k for k, v in METHODS.items()}[m.value]
    return name, off, oo, cont[: nb.value]
