"""CPU-side checks: the C-ABI library loads and exports every symbol include/qsp_recon.h declares,
host-side logic (config mirror, voxel grid, packing), and the loud failure without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "qsp_recon.h")).read()
    return sorted(set(re.findall(r"\b(qsp_[a-z_0-9]+)\s*\(", src)))


@pytest.fixture(scope="module")
def libpath():
    from qsp_slam_b200 import build
    return build.build()


def test_library_exports_every_declared_symbol(libpath):
    lib = ctypes.CDLL(libpath)
    names = _declared()
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), n
    from qsp_slam_b200 import _native
    assert sorted(_native.EXPORTS) == names


def test_no_cpu_fallback(libpath, decoder_state):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from qsp_slam_b200 import _native
    from qsp_slam_b200.decoder import DeviceDecoder
    with pytest.raises(_native.NativeError):
        DeviceDecoder(decoder_state, 0)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "qsp_slam_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt, f


def test_force_key_error_dict_and_configs(tmp_path):
    import json
    from qsp_slam_b200 import synth
    from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict, get_configs
    cfg = synth.optimizer_config("kitti")
    f = tmp_path / "c.json"
    f.write_text(json.dumps(cfg))
    c = get_configs(str(f))
    assert c.optimizer.joint_optim.k4 == 1e7 and c.data_type == "KITTI"
    with pytest.raises(KeyError):
        c.optimizer.nothing
    d = ForceKeyErrorDict(t_cam_obj=None, is_good=False)
    assert d.is_good is False and d["t_cam_obj"] is None


def test_voxel_grid_matches_reference_quirk(golden):
    from qsp_slam_b200.reconstruct.utils import create_voxel_grid
    g = create_voxel_grid(12)
    assert np.array_equal(g, golden["grid12"])
    assert g[1, 1] != g[0, 1]      # fractional y index of the second point (true division)


def test_synth_is_deterministic(decoder_state):
    from qsp_slam_b200 import synth
    st2 = synth.make_decoder_state(0)
    for k in decoder_state:
        assert np.array_equal(decoder_state[k], st2[k])
    pool = np.random.default_rng(0).standard_normal((500, 3)).astype(np.float32) * 0.3
    a = synth.make_object(pool, 5, 100, 32, 16)
    b = synth.make_object(pool, 5, 100, 32, 16)
    assert all(np.array_equal(a[k], b[k]) for k in a)
    assert a["rays"].shape == (48, 3) and np.allclose(a["rays"][:, 2], 1.0) and a["depth"].shape == (32,)


def test_pack_accepts_what_pybind11_delivers():
    """Optimizer._pack (host logic): Fortran-ordered float32 from Eigen, C-ordered float32 / float64 from the Python
    scripts, empty (0,3) and (0,0) members -> one C-ordered float32 array and int32 offsets."""
    from qsp_slam_b200.reconstruct.optimizer import Optimizer
    rng = np.random.default_rng(0)
    arrs = [np.asfortranarray(rng.standard_normal((5, 3)).astype(np.float32)), np.zeros((0, 3), np.float32),
            rng.standard_normal((7, 3)), np.zeros((0, 0), np.float32), rng.standard_normal((4, 3)).astype(np.float32)[::2]]
    cat, off = Optimizer._pack(arrs, 3)
    ref = np.concatenate([np.ascontiguousarray(a, dtype=np.float32).reshape(-1, 3) for a in arrs])
    np.testing.assert_array_equal(cat, ref)
    assert cat.flags.c_contiguous and cat.dtype == np.float32 and off.dtype == np.int32
    assert off.tolist() == [0, 5, 5, 12, 12, 14]
    d, o = Optimizer._pack([np.arange(3.0), np.zeros(0), np.arange(2, dtype=np.float32)], 0)
    assert d.tolist() == [0, 1, 2, 0, 1] and o.tolist() == [0, 3, 3, 5] and d.dtype == np.float32
    e, o = Optimizer._pack([np.zeros((0, 3), np.float32)], 3)
    assert e.shape == (0, 3) and o.tolist() == [0, 0]
