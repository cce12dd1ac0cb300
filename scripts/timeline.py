"""In-situ CUDA-event timeline of one GN iteration of kitti_batch32 (QSP_TIMELINE=1): where the time outside the
two decoder kernels goes.  Prints the library's stderr lines of a warm step."""
import os, sys
os.environ.setdefault("QSP_TIMELINE", "1")      # QSP_TIMELINE=0: the same calls without per-kernel events (ncu passes)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict

st = synth.make_decoder_state(0)
dev = DeviceDecoder(st, 0)
cfg, objs = synth.make_workload(sys.argv[1] if len(sys.argv) > 1 else "kitti_batch32", st)
opt = Optimizer(dev, ForceKeyErrorDict(cfg))
for i in range(3):
    print("---- step", i, file=sys.stderr, flush=True)
    opt.reconstruct_objects(objs)
