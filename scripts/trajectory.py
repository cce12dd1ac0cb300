import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
from oracle import gn_oracle as G
g = dict(np.load('tests/golden/reference_outputs.npz'))
st = synth.make_decoder_state(0); dev = DeviceDecoder(st, 0)
for name, style in (('redwood0', 'redwood'), ('kitti0', 'kitti')):
    p = name + '_'
    cfg = synth.optimizer_config(style, num_depth_samples=50, num_iterations=10)
    for dt in (torch.float32, torch.float64):
        fd = G.FoldedDecoder(st, dt); tr = []
        G.reconstruct_object(fd, G.Params(cfg), g[p+'t_cam_obj'], g[p+'pts'], g[p+'rays'], g[p+'depth'], trace=tr)
        print(name, 'oracle', dt, ' '.join(f"{float(t['loss']):.4f}" for t in tr))
    for prec in ('fp32', 'tensor'):
        dev.set_precision(prec); ls = []
        for k in range(1, 11):
            opt = Optimizer(dev, ForceKeyErrorDict(synth.optimizer_config(style, num_depth_samples=50, num_iterations=k)))
            ls.append(opt.reconstruct_object(g[p+'t_cam_obj'], g[p+'pts'], g[p+'rays'], g[p+'depth']).loss)
        print(name, 'gpu', prec, ' '.join(f"{l:.4f}" for l in ls))
    print(name, 'reference final', float(g[p+'full_loss']))
