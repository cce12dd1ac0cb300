"""Makes ``import reconstruct.optimizer`` / ``reconstruct.utils`` / ``deep_sdf.workspace`` resolve to the
B200 path for the names the C++ core looks up (src/System.cc:118-150, src/LocalMapping.cc:39-69), while
leaving the rest of the reference's ``reconstruct`` package (sequences, detectors) untouched."""
from __future__ import annotations

import importlib
import sys
import types


def install(patch_existing: bool = True) -> None:
    from .reconstruct import optimizer as o, utils as u
    from .deep_sdf import workspace as w

    def _get(name):
        try:
            return importlib.import_module(name)
        except Exception:
            m = types.ModuleType(name)
            sys.modules[name] = m
            return m

    ro, ru, dw = _get("reconstruct.optimizer"), _get("reconstruct.utils"), _get("deep_sdf.workspace")
    ro.Optimizer, ro.MeshExtractor = o.Optimizer, o.MeshExtractor
    ru.ForceKeyErrorDict, ru.get_configs, ru.create_voxel_grid = u.ForceKeyErrorDict, u.get_configs, u.create_voxel_grid
    ru.get_decoder = lambda configs: w.config_decoder(configs.DeepSDF_DIR)   # reconstruct/utils.py:93-95
    dw.config_decoder = w.config_decoder
