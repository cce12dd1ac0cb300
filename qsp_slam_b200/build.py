"""Builds the C-ABI shared library in-tree with nvcc for sm_100a (no JIT cache, no torch extension)."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libqsp_recon.so")
LIB_EXP = os.path.join(HERE, "libqsp_recon_exp.so")     # -DQSP_EXPERIMENTAL: adds the superseded decoder kernels (A/B tests)
SOURCES = [os.path.join(HERE, "csrc", "api.cu")]
HEADERS = [os.path.join(HERE, "csrc", f) for f in ("common.cuh", "decoder_ffma.cuh", "decoder_tc.cuh", "decoder_t3.cuh", "decoder_p3.cuh", "tc_common.cuh", "gn_kernels.cuh", "mesh_kernels.cuh", "mc_table.inc")] + \
          [os.path.join(os.path.dirname(HERE), "include", "qsp_recon.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def needs_build(lib: str = LIB) -> bool:
    if not os.path.isfile(lib):
        return True
    t = os.path.getmtime(lib)
    return any(os.path.getmtime(f) > t for f in SOURCES + HEADERS if os.path.isfile(f))


def build(force: bool = False, verbose: bool = False, experimental: bool = False) -> str:
    """The product library, or (experimental=True) the A/B library that also carries k_decoder_t3, the four-product
    TC_SPLIT mode and the single-CTA variants of k_decoder_tc, selected by QSP_TC_* environment variables."""
    lib = LIB_EXP if experimental else LIB
    if not force and not needs_build(lib):
        return lib
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("QSP_NVCC_EXTRA", "").split()    # experiments only, e.g. -DTC_NG=4
    if experimental:
        extra.append("-DQSP_EXPERIMENTAL")
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + SOURCES + ["-o", lib]
    subprocess.run(cmd, check=True, cwd=os.path.join(HERE, "csrc"))
    return lib


if __name__ == "__main__":
    print(build(force=True, verbose=True))
