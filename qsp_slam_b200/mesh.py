"""Isosurface extraction behind the C ABI (qsp_mesh_*): replaces reconstruct/utils.py:120-141."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N


def marching_cubes(sdf, level: float, spacing: float, origin=(-1.0, -1.0, -1.0)):
    """sdf: (D,D,D) float32 host array or torch CUDA tensor.  Returns (vertices (V,3) f32, faces (F,3) i32).
    vertices = array index * spacing + origin, exactly as the reference maps them (utils.py:136-139)."""
    is_np = isinstance(sdf, np.ndarray)
    if is_np:
        sdf = np.ascontiguousarray(sdf, np.float32)
    D = int(sdf.shape[0])
    assert tuple(sdf.shape) == (D, D, D)
    org = np.asarray(origin, np.float32)
    h = C.c_void_p()
    N.check(N.lib().qsp_mesh_extract(N.ptr(sdf), D, float(level), float(spacing), N.ptr(org),
                                     N.FLAG_HOST_PTRS if is_np else 0, None, C.byref(h)))
    try:
        nv, nf = C.c_int32(), C.c_int32()
        N.check(N.lib().qsp_mesh_sizes(h, C.byref(nv), C.byref(nf)))
        v = np.empty((nv.value, 3), np.float32)
        f = np.empty((nf.value, 3), np.int32)
        N.check(N.lib().qsp_mesh_copy(h, N.ptr(v), N.ptr(f)))
    finally:
        N.lib().qsp_mesh_free(h)
    return v, f
