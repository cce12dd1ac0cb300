"""GPU box: cycle profile of k_solve's stages (experimental build, QSP_SOLVE_PROFILE=1): thread 0 of object 0 prints
clock64() deltas for one warm kitti_single call and one kitti_batch32 call."""
import os, sys
os.environ["QSP_SOLVE_PROFILE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qsp_slam_b200 import synth, build, _native as N
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict

with N.use_library(build.LIB_EXP):
    from qsp_slam_b200.decoder import DeviceDecoder
    st = synth.make_decoder_state(0)
    dev = DeviceDecoder(st, 0, calibrate=False)
    for wl in (sys.argv[1:] or ["kitti_single", "kitti_batch32"]):
        cfg, objs = synth.make_workload(wl, st)
        cfg["optimizer"]["joint_optim"]["num_iterations"] = 2
        opt = Optimizer(dev, ForceKeyErrorDict(cfg))
        print("----", wl, flush=True)
        opt.reconstruct_objects(objs)
        torch.cuda.synchronize()
    dev.close()
