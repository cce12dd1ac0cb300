"""Observed one-iteration errors of the product path against the reference's own outputs (tests/golden) and the oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
from oracle import gn_oracle as G
golden = dict(np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "reference_outputs.npz")))
state = synth.make_decoder_state(0)
dev = DeviceDecoder(state, 0)
fd = G.FoldedDecoder(state)
for precision in ("tensor", "fp32"):
    for name, style in (("kitti0", "kitti"), ("kitti1", "kitti"), ("redwood0", "redwood"), ("redwood1", "redwood")):
        p = name + "_"
        cfg = synth.optimizer_config(style, num_depth_samples=50, num_iterations=1)
        opt = Optimizer(dev, ForceKeyErrorDict(cfg))
        dev.set_precision(precision); opt.set_debug(True)
        out = opt.reconstruct_object(golden[p + "t_cam_obj"], golden[p + "pts"], golden[p + "rays"], golden[p + "depth"], golden[p + "code0"])
        it = opt.get_iteration(0); opt.set_debug(False); dev.set_precision("tensor")
        t_ref, c_ref = golden[p + "it1_t"], golden[p + "it1_code"]
        prm = G.Params(cfg)
        o = G.gn_iteration(fd, prm, torch.inverse(torch.from_numpy(golden[p + "t_cam_obj"])), torch.from_numpy(golden[p + "code0"]),
                           torch.from_numpy(golden[p + "pts"]), torch.from_numpy(golden[p + "rays"]), torch.from_numpy(golden[p + "depth"]))
        rel = lambda a, b, floor=0.0: float(np.abs(a - b).max() / max(np.abs(b).max(), floor))
        print(f"{precision:6s} {name:9s} pose {rel(out.t_cam_obj, t_ref):.1e} code {rel(out.code, c_ref, 0.05):.1e} "
              f"loss {abs(out.loss - float(golden[p + 'it1_loss'])) / abs(float(golden[p + 'it1_loss'])):.1e} | oracle: H {rel(it['H'], o['H'].numpy()):.1e} "
              f"b {rel(it['b'], o['b'].numpy()):.1e} dx {rel(it['dx'], o['dx'].numpy()):.1e}", flush=True)
