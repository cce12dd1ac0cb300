"""The pieces of reconstruct/utils.py the hot path needs (ForceKeyErrorDict :82-84, get_configs :87-90,
create_voxel_grid :98-117), without the addict / skimage / plyfile imports."""
from __future__ import annotations

import json

import numpy as np


class ForceKeyErrorDict(dict):
    """Attribute-style dict that raises KeyError on a missing key (reference: addict.Dict subclass
    whose ``__missing__`` raises, utils.py:82-84).  Nested dicts are wrapped on construction."""

    def __init__(self, *a, **kw):
        super().__init__()
        for k, v in dict(*a, **kw).items():
            self[k] = v

    def __setitem__(self, k, v):
        if isinstance(v, dict) and not isinstance(v, ForceKeyErrorDict):
            v = ForceKeyErrorDict(v)
        super().__setitem__(k, v)

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise KeyError(k)

    def __setattr__(self, k, v):
        self[k] = v

    def __missing__(self, key):
        raise KeyError(key)


def get_configs(cfg_file):
    with open(cfg_file) as f:
        cfg_dict = json.load(f)
    return ForceKeyErrorDict(**cfg_dict)


def create_voxel_grid(vol_dim: int = 128) -> np.ndarray:
    """(vol_dim^3, 3) float32 grid in [-1,1]^3 exactly as the reference builds it, INCLUDING its
    true-division quirk (utils.py:106-107: ``index / vol_dim`` is not floored under torch >= 1.7, so
    the first two columns carry fractional offsets).  The arithmetic is done in float32 like torch's
    ``LongTensor / int -> float32`` promotion."""
    voxel_size = 2.0 / (vol_dim - 1)
    idx = np.arange(vol_dim ** 3, dtype=np.int64)
    v = np.zeros((vol_dim ** 3, 3), np.float32)
    q = idx.astype(np.float32) / np.float32(vol_dim)          # true division, fp32
    v[:, 2] = (idx % vol_dim).astype(np.float32)
    v[:, 1] = np.fmod(q, np.float32(vol_dim))
    v[:, 0] = np.fmod(q / np.float32(vol_dim), np.float32(vol_dim))
    for c in range(3):
        v[:, c] = v[:, c] * np.float32(voxel_size) + np.float32(-1)
    return v


def get_decoder(configs):
    """reconstruct/utils.py:get_decoder - the decoder named by ``configs.DeepSDF_DIR``."""
    from ..deep_sdf.workspace import config_decoder
    return config_decoder(configs.DeepSDF_DIR)


def convert_sdf_voxels_to_mesh(sdf_3d):
    """reconstruct/utils.py:120-141: level-0 isosurface of an (n,n,n) SDF grid (numpy array or torch tensor, host or
    device) with spacing 2/(n-1) and origin (-1,-1,-1).  Returns (vertices (V,3) float32, faces (F,3) int32).  Runs on
    the GPU (qsp_mesh_extract): marching tetrahedra, not scikit-image's Lewiner tables - same level set, different
    triangulation."""
    from ..mesh import marching_cubes
    if hasattr(sdf_3d, "detach") and not sdf_3d.is_cuda:
        sdf_3d = sdf_3d.detach().numpy()
    n = int(sdf_3d.shape[0])
    return marching_cubes(sdf_3d, 0.0, 2.0 / (n - 1), (-1.0, -1.0, -1.0))


def write_mesh_to_ply(v, f, ply_filename_out):
    """reconstruct/utils.py:144-164 (binary little-endian PLY, the layout ``plyfile`` writes)."""
    from ..map_io import write_mesh_to_ply as _w
    _w(v, f, ply_filename_out)
