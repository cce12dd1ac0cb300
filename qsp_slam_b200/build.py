"""Builds the C-ABI shared library in-tree with nvcc for sm_100a (no JIT cache, no torch extension)."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libqsp_recon.so")
SOURCES = [os.path.join(HERE, "csrc", "api.cu")]
HEADERS = [os.path.join(HERE, "csrc", f) for f in ("common.cuh", "decoder_ffma.cuh", "decoder_tc.cuh", "decoder_t3.cuh", "decoder_p3.cuh", "tc_common.cuh", "gn_kernels.cuh", "mesh_kernels.cuh")] + \
          [os.path.join(os.path.dirname(HERE), "include", "qsp_recon.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def needs_build() -> bool:
    if not os.path.isfile(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(f) > t for f in SOURCES + HEADERS if os.path.isfile(f))


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("QSP_NVCC_EXTRA", "").split()    # experiments only, e.g. -DTC_NG=4
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + SOURCES + ["-o", LIB]
    subprocess.run(cmd, check=True, cwd=os.path.join(HERE, "csrc"))
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
