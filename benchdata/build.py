"""Build the workload generator: libsynth_gpu.so (nvcc, sm_100a) and libsynth_host.so (g++)."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
GPU_LIB = os.path.join(HERE, "libsynth_gpu.so")
HOST_LIB = os.path.join(HERE, "libsynth_host.so")


def _stale(lib: str, deps: list[str]) -> bool:
    return not os.path.exists(lib) or any(os.path.getmtime(d) > os.path.getmtime(lib) for d in deps)


def build(force: bool = False) -> None:
    core = os.path.join(HERE, "synth_core.h")
    src = os.path.join(HERE, "synth_host.cpp")
    if force or _stale(HOST_LIB, [core, src]):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", HOST_LIB, src], check=True)
    src = os.path.join(HERE, "synth_gpu.cu")
    if force or _stale(GPU_LIB, [core, src]):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
                        "-Xcompiler", "-fPIC", "-shared", "-o", GPU_LIB, src], check=True)


if __name__ == "__main__":
    build(force=True)
