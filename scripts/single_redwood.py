"""One Redwood-style object (BASELINE config 3: 4k rays x 128 samples) through Optimizer.reconstruct_object."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
state = synth.make_decoder_state(0)
dev = DeviceDecoder(state, 0)
cfg, objs = synth.make_workload("redwood_chair", state, decode=lambda c, p: dev.decode_sdf(c, p))
o = objs[0]
opt = Optimizer(dev, ForceKeyErrorDict(cfg))
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    t0 = time.perf_counter()
    r = opt.reconstruct_object(o["t_cam_obj"], o["pts"], o["rays"], o["depth"])
    print(f"call {i}: {1e3 * (time.perf_counter() - t0):.3f} ms is_good={r.is_good}", flush=True)
