"""Experiment: one 32-object step as two 16-object half batches on two streams (two handles = two workspaces), so that
one half's small kernels and decoder tails overlap the other half's decoder kernels."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
state = synth.make_decoder_state(0)
devs = [DeviceDecoder(state, 0), DeviceDecoder(state, 0)]
cfg, objs = synth.make_workload("kitti_batch32", state, decode=lambda c, p: devs[0].decode_sdf(c, p))
opts = [Optimizer(d, ForceKeyErrorDict(cfg)) for d in devs]
def dev_pack(opt, group):
    pk = opt.pack(group)
    return {k: (torch.from_numpy(v).cuda() if k in ("T", "pts", "rays", "depth") else v) for k, v in pk.items()}
full = dev_pack(opts[0], objs); rec_full = torch.empty((32, 82), device="cuda")
for nsplit in (1, 2, 4):
    groups = [objs[i * 32 // nsplit:(i + 1) * 32 // nsplit] for i in range(nsplit)]
    hs = [DeviceDecoder(state, 0) for _ in range(nsplit)] if nsplit > 2 else devs[:nsplit]
    os_ = [Optimizer(d, ForceKeyErrorDict(cfg)) for d in hs]
    pks = [dev_pack(os_[i], groups[i]) for i in range(nsplit)]
    recs = [torch.empty((len(g), 82), device="cuda") for g in groups]
    streams = [torch.cuda.Stream() for _ in range(nsplit)]
    def step():
        for i in range(nsplit):
            os_[i].reconstruct_packed_device(pks[i], recs[i], sync=False, stream=streams[i].cuda_stream)
    for _ in range(3): step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5): step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    good = sum(int((r[:, 80] != 0).sum().item()) for r in recs)
    print(f"{nsplit} stream(s): {dt * 1e3:.2f} ms per 32-object step, {32 / dt:.0f} objects/s, good {good}", flush=True)
