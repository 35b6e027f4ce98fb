"""The C boundary: include/press_gpu.h compiles as C99, libpress_gpu.so exports every function the
header declares, and (on a B200) tests/c_abi/dropin.c -- the test_<method>_gpu functions a
maintainer adds to press/test.c -- round trips the three fixture reads with the reference's bytes."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "press_gpu.h")
CABI = os.path.join(ROOT, "tests", "c_abi")


def _declared_functions():
    src = open(HDR).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"typedef struct \w+ \{.*?\}\s*\w+;", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(\w+)\s*\([^;{}]*\)\s*;", src)))


def _build_lib():
    from honours_b200 import build
    return build.build()


def test_header_symbols_are_exported():
    names = _declared_functions()
    assert len(names) > 50 and "press_gpu_batch_press" in names and "svb12_zd_gpu_press" in names
    L = C.CDLL(_build_lib())
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_library_has_no_generator_or_debug_exports():
    """the product ABI must not carry the bench workload generator or debug hooks."""
    out = subprocess.run(["nm", "-D", "--defined-only", _build_lib()], capture_output=True, text=True, check=True).stdout
    assert "synth" not in out and "debug_aux" not in out


def test_dropin_compiles_as_c99_and_links():
    _build_lib()
    subprocess.run(["make", "-s", "-B", "-C", CABI], check=True)
    assert os.path.exists(os.path.join(CABI, "dropin"))


@pytest.mark.gpu
def test_dropin_round_trips_fixture_reads(oracle, tmp_path):
    """config 1 through compiled C: every TSV line (method, n, bytes, FNV-1a-64) must equal the
    oracle's stream for that read; the Huffman table goes in as SymbolEncoder* / huffman_node*."""
    from oracle import cpu_impl
    _build_lib()
    subprocess.run(["make", "-s", "-C", CABI], check=True)
    z = np.load(os.path.join(ROOT, "tests", "golden", "three_reads.npz"))
    samples, offsets = z["samples"].astype(np.int16), z["offsets"].astype(np.uint64)
    # plus two short synthetic reads (a q > 0 read among them)
    rng = np.random.default_rng(5)
    extra = [((100 + np.cumsum(rng.integers(-6, 7, 5000))) << 2).astype(np.int16),
             (475 + np.cumsum(rng.integers(-25, 26, 40))).astype(np.int16)]
    reads = [samples[int(offsets[r]):int(offsets[r + 1])] for r in range(3)] + extra
    off = np.concatenate([[0], np.cumsum([r.size for r in reads])]).astype(np.uint64)
    p = tmp_path / "reads.bin"
    with open(p, "wb") as f:
        f.write(np.uint64(len(reads)).tobytes())
        f.write(off.tobytes())
        f.write(np.concatenate(reads).tobytes())
    table = os.path.join(ROOT, "honours_b200", "data", "NA12878_zd.huffman")
    r = subprocess.run([os.path.join(CABI, "dropin"), str(p), table], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    name = {"svb12_zd_gpu": "svb16_zd", "svb12_zd_nosimd_gpu": "svb16_zd", "vbe21_zd_gpu": "vbe21_zd",
            "hasgam_vbsse21_zdq_gpu": "ex_zd", "shuffman_vbe21_zd_gpu": "shuff_vbe21_zd",
            "svb0124_zd_gpu": "svb0124_zd", "uint_zd_gpu": "uint_zd"}
    lines = [ln.split("\t") for ln in r.stdout.strip().split("\n")]
    assert len(lines) == 7 * len(reads)
    for i, (m, n, nbytes, h) in enumerate(lines):
        s = reads[i // 7]
        rc, st = oracle.press(name[m], s)
        assert rc == 0 and int(n) == s.size and int(nbytes) == len(st), (m, n)
        assert h == f"{cpu_impl.fnv1a64(st):016x}", (m, n)
