"""Out-of-bounds check of the CUDA path without compute-sanitizer (closed on the GPU pool): with QSP_GUARD=1 the
library allocates every workspace buffer at exactly the size a call asks for, between two 256-byte canary bands
(`qsp_check_guards`, include/qsp_recon.h).  Each case below runs on a FRESH handle, so its buffers are sized for that
case alone: a kernel that writes one element past a ragged tail lands in a band."""
import numpy as np
import pytest
import torch

from qsp_slam_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture()
def guarded(decoder_state, monkeypatch):
    from qsp_slam_b200.decoder import DeviceDecoder
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    monkeypatch.setenv("QSP_GUARD", "1")
    d = DeviceDecoder(decoder_state, 0)
    yield d
    d.close()


def make_opt(dev, style, iters, samples):
    from qsp_slam_b200.reconstruct.optimizer import Optimizer
    from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
    return Optimizer(dev, ForceKeyErrorDict(synth.optimizer_config(style, num_depth_samples=samples, num_iterations=iters)))


def test_guard_mode_is_off_by_default(decoder_state, guarded):
    """An intact guarded handle reports 0; a handle created without QSP_GUARD holds no guarded buffer (-1)."""
    import os
    from qsp_slam_b200.decoder import DeviceDecoder
    x = np.zeros((5, 3), np.float32)
    guarded.decode_sdf(np.zeros(64, np.float32), x)
    assert guarded.check_guards() == 0
    os.environ.pop("QSP_GUARD")
    try:
        plain = DeviceDecoder(decoder_state, 0)
        plain.decode_sdf(np.zeros(64, np.float32), x)
        assert plain.check_guards() == -1
        plain.close()
    finally:
        os.environ["QSP_GUARD"] = "1"


def test_a_write_past_a_buffer_is_reported(guarded):
    """Positive control: one byte written just past a guarded buffer makes the check fail and name the buffer."""
    from qsp_slam_b200 import _native as N
    guarded.decode_sdf(np.zeros(64, np.float32), np.zeros((5, 3), np.float32))
    assert guarded.check_guards() == 0
    N.check(N.lib().qsp_guard_selftest(guarded.handle))
    with pytest.raises(N.NativeError, match=r"guard bands overwritten: \w+ \(above\)"):
        guarded.check_guards()


@pytest.mark.parametrize("precision", ["tensor", "fp32"])
@pytest.mark.parametrize("samples", [7, 33, 100])
def test_reconstruction_stays_inside_its_buffers(golden, guarded, precision, samples):
    guarded.set_precision(precision)
    opt = make_opt(guarded, "redwood", 2, samples)
    p = "redwood0_"
    far = golden[p + "rays"].copy()
    far[:, 0] += 5.0
    e3, e1 = np.zeros((0, 3), np.float32), np.zeros(0, np.float32)
    dets = [dict(t_cam_obj=golden[p + "t_cam_obj"], pts=golden[p + "pts"][:101], rays=golden[p + "rays"], depth=golden[p + "depth"]),
            dict(t_cam_obj=golden[p + "t_cam_obj"], pts=golden[p + "pts"], rays=far, depth=golden[p + "depth"]),
            dict(t_cam_obj=golden[p + "t_cam_obj"], pts=golden[p + "pts"][:1], rays=e3, depth=e1),
            dict(t_cam_obj=golden["redwood1_t_cam_obj"], pts=golden["redwood1_pts"], rays=golden["redwood1_rays"][:41],
                 depth=golden["redwood1_depth"][:37])]
    res = opt.reconstruct_objects(dets)
    assert [r.is_good for r in res] == [True, False, False, True]
    assert guarded.check_guards() == 0


def test_kitti_single_object_and_pose_only_stay_inside_their_buffers(golden, guarded):
    opt = make_opt(guarded, "kitti", 3, 50)
    p = "kitti0_"
    assert opt.reconstruct_object(golden[p + "t_cam_obj"], golden[p + "pts"], golden[p + "rays"], golden[p + "depth"]).is_good
    assert guarded.check_guards() == 0
    opt.num_iterations_pose_only = 5
    opt.estimate_pose_cam_obj(golden[p + "po_in"].copy(), float(golden[p + "po_scale"]), golden[p + "pts"][:777], golden[p + "code0"])
    assert guarded.check_guards() == 0


@pytest.mark.parametrize("n", [1, 63, 64, 65, 127, 129, 9473, 18945])
def test_decoder_entry_points_stay_inside_their_buffers(decoder_state, monkeypatch, n):
    """Row counts around the tile sizes (64 / 128 rows per CTA, 148 CTAs)."""
    from qsp_slam_b200.decoder import DeviceDecoder
    monkeypatch.setenv("QSP_GUARD", "1")
    d = DeviceDecoder(decoder_state, 0)
    rng = np.random.default_rng(n)
    x = (rng.standard_normal((n, 3)) * 0.4).astype(np.float32)
    z = (rng.standard_normal(64) * 0.1).astype(np.float32)
    for mode in ("tensor", "fast", "fp32"):
        d.decode_sdf(z, x, mode=mode)
    for mode in ("tensor", "fp32"):
        d.decode_sdf_jacobian(z, x, mode=mode)
    assert d.check_guards() == 0
    d.close()


@pytest.mark.parametrize("D", [17, 48])
def test_mesh_extraction_stays_inside_its_buffers(guarded, D):
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    ex = MeshExtractor(guarded, voxels_dim=D)
    code = (np.random.default_rng(D).standard_normal(64) * 0.1).astype(np.float32)
    ex.extract_mesh_from_code(code)
    ex.decode_grid(code[None])
    assert guarded.check_guards() == 0
