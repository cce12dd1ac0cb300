"""Where the end-to-end time of a 32-object call goes beyond the device-resident step: host packing, the C call with
host pointers (h2d + loop + d2h), the same call with device pointers, result unpacking.  Wall clock, medians of 10."""
import os, sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth, _native as N
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict

state = synth.make_decoder_state(0)
dev = DeviceDecoder(state, 0)
cfg, objs = synth.make_workload("kitti_batch32", state, objects=32, pool_size=4096)
opt = Optimizer(dev, ForceKeyErrorDict(cfg))
def med(f, n=10):
    f(); f()
    ts = []
    for _ in range(n):
        torch.cuda.synchronize(); t = time.perf_counter(); f(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t)
    return 1e3 * float(np.median(ts))
print("reconstruct_objects (public API)      %.3f ms" % med(lambda: opt.reconstruct_objects(objs)))
print("reconstruct_objects_records           %.3f ms" % med(lambda: opt.reconstruct_objects_records(objs)))
print("pack only                             %.3f ms" % med(lambda: opt.pack(objs)))
pk = opt.pack(objs)
rec = np.empty((32, N.RECORD_LEN), np.float32)
prm = opt._params()
def host_call():
    N.check(N.lib().qsp_reconstruct_batch(dev.handle, C.byref(prm), 32, N.ptr(pk["T"]), None, N.ptr(pk["pts"]), N.i32p(pk["pts_off"]),
                                          N.ptr(pk["rays"]), N.i32p(pk["rays_off"]), N.ptr(pk["depth"]), N.i32p(pk["depth_off"]),
                                          N.ptr(rec), N.FLAG_HOST_PTRS, None))
print("C call, packed host buffers           %.3f ms" % med(host_call))
dpk = {k: (torch.from_numpy(v).cuda() if k in ("T", "pts", "rays", "depth") else v) for k, v in pk.items()}
drec = torch.empty((32, N.RECORD_LEN), device="cuda")
print("C call, device buffers + sync         %.3f ms" % med(lambda: opt.reconstruct_packed_device(dpk, drec)))
print("unpack 32 results                     %.3f ms" % med(lambda: [opt._result(rec[i]) for i in range(32)]))
