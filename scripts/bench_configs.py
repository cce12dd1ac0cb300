"""Timings of the other BASELINE.json configs (parity-test cases, not the bench line): single KITTI car,
Redwood chair (1 object and 4 yaw hypotheses), 128^3 grid decode x 64 codes, pose-only GN.
usage: python scripts/bench_configs.py > profiles/rXX_configs.jsonl"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer, MeshExtractor
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict, create_voxel_grid

state = synth.make_decoder_state(0)
dev = DeviceDecoder(state, 0)
decode = lambda c, p: dev.decode_sdf(c, p)

def timed(fn, reps=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    return float(np.median(ts))

out = []
for name, kw in (("kitti_single", {}), ("redwood_chair", {})):
    cfg, objs = synth.make_workload(name, state, decode=decode)
    opt = Optimizer(dev, ForceKeyErrorDict(cfg)); o = objs[0]
    t = timed(lambda: opt.reconstruct_object(o["t_cam_obj"], o["pts"], o["rays"], o["depth"]))
    opt.set_debug(True); r = opt.reconstruct_object(o["t_cam_obj"], o["pts"], o["rays"], o["depth"]); it = opt.get_iteration(5); opt.set_debug(False)
    out.append(dict(config=name, objects=1, seconds=t, objects_per_s=1 / t, is_good=r.is_good, loss=r.loss,
                    in_ball=it["n_inball"], candidates=it["n_cand"], kept_band=it["n_keep"], api="Optimizer.reconstruct_object (host buffers)"))
    if name == "redwood_chair":
        t4 = timed(lambda: opt.reconstruct_object_hypotheses(o["t_cam_obj"], o["pts"], o["rays"], o["depth"], flip_sample_num=4))
        out.append(dict(config=name + "_4_hypotheses", objects=4, seconds=t4, objects_per_s=4 / t4, api="reconstruct_object_hypotheses"))
# pose-only, 32 objects x 2000 points
cfg, objs = synth.make_workload("kitti_batch32", state, decode=decode)
opt = Optimizer(dev, ForceKeyErrorDict(cfg))
Ts, ss = [], []
for o in objs:
    T = o["t_cam_obj"].copy(); s = np.cbrt(np.linalg.det(T[:3, :3].astype(np.float64))); T[:3, :3] /= s; Ts.append(T); ss.append(float(s))
z = [np.zeros(64, np.float32)] * len(objs)
tp = timed(lambda: opt.estimate_poses_cam_obj(Ts, ss, [o["pts"] for o in objs], z))
out.append(dict(config="pose_only_32_objects", objects=32, seconds=tp, objects_per_s=32 / tp, iterations=5))
# grid decode: 64 codes x 128^3 (device resident), exact (fp16-pair) and single-pass fp16
grid = torch.from_numpy(create_voxel_grid(128)).cuda()
codes = (torch.randn(64, 64, generator=torch.Generator().manual_seed(0)) * 0.1).cuda()
res = torch.empty((64, grid.shape[0]), device="cuda")
for mode in ("tensor", "fast"):
    for _ in range(2): dev.decode_sdf_device(codes, grid, res, mode=mode)
    ms = dev.last_kernel_ms(); n = 64 * grid.shape[0]
    out.append(dict(config="mesh_grid_128cubed_x64", mode=mode, seconds=ms / 1e3, evals_per_s=n / (ms / 1e3),
                    tflops_algorithmic=n * 3.67104e6 / (ms * 1e-3) / 1e12))
# the whole of MeshExtractor.extract_mesh_from_code (host code in, host vertices/faces out), 64 codes
from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
for D in (64, 128):
    mx = MeshExtractor(dev, 64, D)
    cz = codes.cpu().numpy()
    mx.extract_mesh_from_code(cz[0])
    tm = timed(lambda: [mx.extract_mesh_from_code(c) for c in cz])
    nv = sum(mx.extract_mesh_from_code(c).vertices.shape[0] for c in cz[:4]) / 4
    out.append(dict(config="mesh_extract_%dcubed_x64" % D, seconds=tm, meshes_per_s=64 / tm, ms_per_mesh=tm / 64 * 1e3,
                    mean_vertices=nv, api="MeshExtractor.extract_mesh_from_code (host code in, host mesh out)"))
for o in out: print(json.dumps(o))
