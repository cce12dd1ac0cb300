"""Latency of the calls the C++ core makes - one object per call (src/LocalMapping_util.cc:213,244-245,705-753) - under
the current environment (QSP_FUSED / QSP_CONCURRENT A/B switches): median of 9 calls after 3 warm-ups, host buffers in,
host result out.  Prints one JSON line."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict

state = synth.make_decoder_state(0)
dev = DeviceDecoder(state, 0)

def timed(fn, reps=9, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return 1e3 * float(np.median(ts))

out = {"env": {k: os.environ.get(k) for k in ("QSP_FUSED", "QSP_CONCURRENT", "QSP_PDL") if os.environ.get(k) is not None}}
for name in ("kitti_single", "redwood_chair"):
    cfg, objs = synth.make_workload(name, state)
    opt = Optimizer(dev, ForceKeyErrorDict(cfg)); o = objs[0]
    out[name + "_ms"] = timed(lambda: opt.reconstruct_object(o["t_cam_obj"], o["pts"], o["rays"], o["depth"]))
    r = opt.reconstruct_object(o["t_cam_obj"], o["pts"], o["rays"], o["depth"])
    out[name + "_loss"] = r.loss
    if name == "redwood_chair":
        out["hypotheses4_ms"] = timed(lambda: opt.reconstruct_object_hypotheses(o["t_cam_obj"], o["pts"], o["rays"], o["depth"], flip_sample_num=4), reps=5)
print(json.dumps(out))
