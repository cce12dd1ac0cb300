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
    return _take_mesh(h)


def _take_mesh(h):
    """Copies a qsp_mesh to host arrays and frees it."""
    try:
        nv, nf = C.c_int32(), C.c_int32()
        N.check(N.lib().qsp_mesh_sizes(h, C.byref(nv), C.byref(nf)))
        v = np.empty((nv.value, 3), np.float32)
        f = np.empty((nf.value, 3), np.int32)
        N.check(N.lib().qsp_mesh_copy(h, N.ptr(v), N.ptr(f)))
    finally:
        N.lib().qsp_mesh_free(h)
    return v, f


def extract_mesh(decoder_handle, code, level: float, spacing: float, origin=(-1.0, -1.0, -1.0), exact_grid: bool = False,
                 return_grid_dim: int = 0):
    """Grid decode + isosurface on the device over the handle's voxel grid (qsp_set_voxel_grid): optimizer.py:292-304.
    ``exact_grid``: fp32-equivalent SDF at every grid point instead of only where the isosurface depends on it.
    ``return_grid_dim`` = D also returns the (D,D,D) SDF grid that was triangulated."""
    code = np.ascontiguousarray(code, np.float32).reshape(-1)[:64]
    org = np.asarray(origin, np.float32)
    grid = np.empty((return_grid_dim,) * 3, np.float32) if return_grid_dim else None
    h = C.c_void_p()
    N.check(N.lib().qsp_extract_mesh(decoder_handle, N.ptr(code), float(level), float(spacing), N.ptr(org),
                                     N.ptr(grid) if grid is not None else None,
                                     N.FLAG_HOST_PTRS | (N.FLAG_EXACT_GRID if exact_grid else 0), None, C.byref(h)))
    v, f = _take_mesh(h)
    return (v, f, grid) if grid is not None else (v, f)
