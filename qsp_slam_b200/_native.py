"""ctypes binding of include/qsp_recon.h.  There is no CPU fallback: if the shared library is missing
or no B200 is present, calls raise."""
from __future__ import annotations

import contextlib
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libqsp_recon.so")

FLAG_HOST_PTRS = 1
FLAG_SYNC = 2
FLAG_FAST_FP16 = 4
FLAG_FP32_CORES = 8
PREC_TENSOR = 0
PREC_FP32_CORES = 1
FLAG_EXACT_GRID = 16
FLAG_MESH_TETS = 32
MODE_FLAGS = {"tensor": 0, "fast": FLAG_FAST_FP16, "fp32": FLAG_FP32_CORES}
CODE_LEN = 64
JAC_STRIDE = 68
RECORD_LEN = 82
NOALLOC_MAX_N = 4194304                    # QSP_NOALLOC_MAX_N of include/qsp_recon.h

EXPORTS = ["qsp_decoder_create", "qsp_decoder_destroy", "qsp_decode_sdf", "qsp_decode_sdf_jacobian",
           "qsp_reconstruct_batch", "qsp_get_call_stats", "qsp_estimate_pose_batch", "qsp_set_precision", "qsp_set_first_pass_margin", "qsp_check_guards", "qsp_guard_selftest", "qsp_set_debug", "qsp_get_iteration",
           "qsp_get_band_indices", "qsp_get_loss_rows", "qsp_rotation_prior", "qsp_exp_map", "qsp_robust_res", "qsp_robust_res_noalloc", "qsp_sdf_to_occupancy", "qsp_point_pose_jacobian", "qsp_mesh_extract", "qsp_set_voxel_grid", "qsp_extract_mesh", "qsp_mesh_sizes", "qsp_mesh_copy", "qsp_mesh_free", "qsp_last_error", "qsp_launch_count", "qsp_kernel_ms", "qsp_version"]

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)


class DecoderWeights(C.Structure):
    _fields_ = [("weight_v", _f32p * 8), ("weight_g", _f32p * 8), ("bias", _f32p * 8),
                ("lin8_weight", _f32p), ("lin8_bias", _f32p)]


class GnParams(C.Structure):
    _fields_ = [("k1", C.c_float), ("k2", C.c_float), ("k3", C.c_float), ("k4", C.c_float),
                ("b1", C.c_float), ("b2", C.c_float), ("lr", C.c_float), ("s_damp", C.c_float),
                ("cut_off", C.c_float), ("num_iterations", C.c_int32), ("num_depth_samples", C.c_int32),
                ("reserved", C.c_int32)]


class NativeError(RuntimeError):
    pass


_lib = None


@contextlib.contextmanager
def use_library(path: str):
    """Tests only: bind the package to another build of the library (``libqsp_recon_exp.so``, which also carries the
    superseded decoder kernels) for the duration of the block.  Handles created inside must be closed inside."""
    global _lib
    old = _lib
    _lib = _load(path)
    try:
        yield _lib
    finally:
        _lib = old


def lib():
    global _lib
    if _lib is None:
        _lib = _load(LIB_PATH)
    return _lib


def _load(path: str):
    if not os.path.isfile(path):
        raise NativeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(path)
    vp = C.c_void_p
    L.qsp_decoder_create.argtypes = [C.POINTER(DecoderWeights), C.c_int, C.POINTER(vp)]
    L.qsp_decoder_destroy.argtypes = [vp]
    L.qsp_decoder_destroy.restype = None
    L.qsp_decode_sdf.argtypes = [vp, vp, C.c_int32, vp, C.c_int64, vp, C.c_uint32, vp]
    L.qsp_decode_sdf_jacobian.argtypes = [vp, vp, vp, C.c_int64, vp, vp, C.c_uint32, vp]
    L.qsp_reconstruct_batch.argtypes = [vp, C.POINTER(GnParams), C.c_int32, vp, vp, vp, _i32p, vp, _i32p,
                                        vp, _i32p, vp, C.c_uint32, vp]
    L.qsp_estimate_pose_batch.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp, vp, _i32p, vp, C.c_uint32, vp]
    L.qsp_get_call_stats.argtypes = [vp, _i32p, _f32p, _i32p]
    L.qsp_set_debug.argtypes = [vp, C.c_int32]
    L.qsp_set_precision.argtypes = [vp, C.c_int32]
    L.qsp_set_first_pass_margin.argtypes = [vp, C.c_float, C.c_float]
    L.qsp_check_guards.argtypes = [vp]
    L.qsp_guard_selftest.argtypes = [vp]
    L.qsp_get_iteration.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, vp, vp]
    L.qsp_get_band_indices.argtypes = [vp, C.c_int32, C.c_int32, vp, vp, C.c_int32]
    L.qsp_get_loss_rows.argtypes = [vp, C.c_int32, C.c_int32, C.c_int32, vp, C.c_int32]
    L.qsp_rotation_prior.argtypes = [C.c_int32, vp, vp]
    L.qsp_exp_map.argtypes = [C.c_int32, vp, C.c_int32, C.c_int32, vp]
    L.qsp_robust_res.argtypes = [C.c_int32, vp, C.c_int32, C.c_float, vp, vp, vp]
    L.qsp_robust_res_noalloc.argtypes = [C.c_int32, vp, C.c_int32, C.c_float, vp, vp, vp, vp]
    L.qsp_sdf_to_occupancy.argtypes = [C.c_int32, vp, C.c_int64, C.c_float, vp]
    L.qsp_point_pose_jacobian.argtypes = [C.c_int32, vp, C.c_int32, C.c_int32, vp]
    L.qsp_mesh_extract.argtypes = [vp, C.c_int32, C.c_float, C.c_float, vp, C.c_uint32, vp, C.POINTER(vp)]
    L.qsp_set_voxel_grid.argtypes = [vp, vp, C.c_int32, C.c_uint32, vp]
    L.qsp_extract_mesh.argtypes = [vp, vp, C.c_float, C.c_float, vp, vp, C.c_uint32, vp, C.POINTER(vp)]
    L.qsp_mesh_sizes.argtypes = [vp, _i32p, _i32p]
    L.qsp_mesh_copy.argtypes = [vp, vp, vp]
    L.qsp_mesh_free.argtypes = [vp]
    L.qsp_mesh_free.restype = None
    L.qsp_last_error.restype = C.c_char_p
    L.qsp_launch_count.restype = C.c_int64
    L.qsp_kernel_ms.argtypes = [vp, _f32p]
    L.qsp_version.restype = C.c_char_p
    return L


def check(status: int):
    if status < 0:
        raise NativeError(f"qsp_recon status {status}: {lib().qsp_last_error().decode()}")
    return status


def f32(a) -> np.ndarray:
    """C-contiguous float32 view/copy (inputs from pybind11/eigen.h arrive Fortran-ordered)."""
    return np.ascontiguousarray(a, dtype=np.float32)


def ptr(a):
    """void* of a numpy array, a torch tensor (host or device) or None."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(a.data_ptr())


def i32p(a: np.ndarray):
    return a.ctypes.data_as(_i32p)


def launch_count() -> int:
    return int(lib().qsp_launch_count())
