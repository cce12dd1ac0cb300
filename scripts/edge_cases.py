"""Edge inputs through the public API on the GPU: empty surface points, empty rays, K = 0 depths, tiny inputs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
state = synth.make_decoder_state(0)
dev = DeviceDecoder(state, 0)
cfg, objs = synth.make_workload("kitti_single", state, pool_size=2048, n_surface=64, n_fg=24, n_bg=8, samples=20, iters=2)
o = objs[0]
opt = Optimizer(dev, ForceKeyErrorDict(cfg))
e3 = np.zeros((0, 3), np.float32)
for name, args in (("normal", (o["t_cam_obj"], o["pts"], o["rays"], o["depth"])),
                   ("no surface points", (o["t_cam_obj"], e3, o["rays"], o["depth"])),
                   ("no rays", (o["t_cam_obj"], o["pts"], e3, np.zeros(0, np.float32))),
                   ("no depths (all background)", (o["t_cam_obj"], o["pts"], o["rays"], np.zeros(0, np.float32))),
                   ("one point one ray", (o["t_cam_obj"], o["pts"][:1], o["rays"][:1], o["depth"][:1]))):
    try:
        r = opt.reconstruct_object(*args)
        print(f"{name:28s} is_good={r.is_good} loss={r.loss} t_cam_obj={'None' if r.t_cam_obj is None else r.t_cam_obj.shape}")
    except Exception as ex:
        print(f"{name:28s} RAISED {type(ex).__name__}: {ex}")
print("decode_sdf empty:", dev.decode_sdf(np.zeros(64, np.float32), e3).shape)
