"""End-to-end time of MeshExtractor.extract_mesh_from_code (host code in, host mesh out) and its parts."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from qsp_slam_b200 import synth
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
from qsp_slam_b200.mesh import marching_cubes

st = synth.make_decoder_state(0)
dev = DeviceDecoder(st, 0)
rng = np.random.default_rng(3)
def timed(f, n=5):
    f(); torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(n): r = f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / n, r
for D in (64, 128):
    ex = MeshExtractor(dev, 64, D)
    code = (rng.standard_normal(64) * 0.1).astype(np.float32)
    t_all, mesh = timed(lambda: ex.extract_mesh_from_code(code))
    t_dec, grid = timed(lambda: ex.decode_grid(code))
    t_mc, _ = timed(lambda: marching_cubes(grid, 0.0, 2.0 / (D - 1)))
    out = dict(D=D, extract_ms=t_all * 1e3, decode_grid_host_ms=t_dec * 1e3, marching_host_ms=t_mc * 1e3,
               vertices=int(mesh.vertices.shape[0]), faces=int(mesh.faces.shape[0]))
    if hasattr(ex, "extract_mesh_from_code_reference_grid"):
        t_ref, mref = timed(lambda: ex.extract_mesh_from_code_reference_grid(code))
        out["extract_full_precise_ms"] = t_ref * 1e3
        out["identical"] = bool(np.array_equal(mref.vertices, mesh.vertices) and np.array_equal(mref.faces, mesh.faces))
    print(json.dumps(out), flush=True)
