"""TEST INFRASTRUCTURE ONLY - generates tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python -m oracle.make_golden
The fixtures are small (a few hundred KB) and are the pins for ``oracle/gn_oracle.py`` and,
through it, for the CUDA path.  Configuration of record: fp32, ``torch.set_num_threads(1)``.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from qsp_slam_b200 import synth  # noqa: E402
from oracle import ref_harness as rh  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
DECODER_SEED = 0


def small_workload(state, style):
    over = dict(n_surface=300, n_fg=64, n_bg=32, samples=50, objects=2)
    name = "kitti_single" if style == "kitti" else "redwood_chair"
    return synth.make_workload(name, state, decoder_seed=DECODER_SEED, pool_size=4096, **over)


def main():
    torch.set_num_threads(1)
    ref = rh.load()
    state = synth.make_decoder_state(DECODER_SEED)
    dec = rh.build_decoder(state)
    os.makedirs(OUT, exist_ok=True)
    g = {}

    # --- decoder: SDF and input Jacobian (loss_utils.py:51-103) -------------------------------
    rng = np.random.default_rng(7)
    x = synth._ball_points(rng, 384)
    z = (rng.standard_normal(64) * 0.05).astype(np.float32)
    with torch.no_grad():
        g["dec_x"], g["dec_z"] = x, z
        g["dec_sdf"] = ref.loss_utils.decode_sdf(dec, torch.from_numpy(z), torch.from_numpy(x)).numpy()
    y, jac = ref.loss_utils.get_batch_sdf_jacobian(dec, torch.from_numpy(z), torch.from_numpy(x), 1)
    g["dec_jac_sdf"], g["dec_jac"] = y.reshape(-1).numpy(), jac.reshape(-1, 67).numpy()

    # --- voxel grid and grid decode (utils.py:98-117, optimizer.py:296) -----------------------
    grid = ref.utils.create_voxel_grid(vol_dim=12)
    g["grid12"] = grid.numpy()
    with torch.no_grad():
        g["grid12_sdf"] = ref.loss_utils.decode_sdf(dec, torch.from_numpy(z), grid).numpy()

    # --- small helpers ---------------------------------------------------------------------------
    xs = np.array([[0.1, -0.2, 0.3, 0.05, -0.02, 0.04, 0.03],
                   [0.1, -0.2, 0.3, 0.05, -0.02, 0.04, -0.03],
                   [0.1, -0.2, 0.3, 0.0, 0.0, 0.0, 0.02],
                   [0.1, -0.2, 0.3, 0.0, 0.0, 0.0, 0.0],
                   [1e-3, 2e-3, -1e-3, 1e-4, -2e-4, 3e-4, 1e-5]], np.float32)
    g["exp_in"] = xs
    g["exp_sim3"] = np.stack([ref.loss_utils.exp_sim3(torch.from_numpy(v)).numpy() for v in xs])
    g["exp_se3"] = np.stack([ref.loss_utils.exp_se3(torch.from_numpy(v[:6])).numpy() for v in xs])
    r = (rng.standard_normal(200) * 0.05).astype(np.float32)
    r[:3] = 0.0
    rr, loss, w = ref.loss_utils.get_robust_res(torch.from_numpy(r.copy()), 0.025)
    g["rob_in"], g["rob_res"], g["rob_loss"], g["rob_w"] = r, rr.reshape(-1).numpy(), loss.numpy(), w.reshape(-1).numpy()

    # --- per-term and per-iteration outputs on two small objects per style -----------------------
    for style in ("kitti", "redwood"):
        cfg, objs = small_workload(state, style)
        opt1 = ref.optimizer.Optimizer(dec, rh.make_configs(synth.optimizer_config(
            style, num_depth_samples=50, num_iterations=1)))
        optN = ref.optimizer.Optimizer(dec, rh.make_configs(cfg))
        for i, o in enumerate(objs):
            p = f"{style}{i}_"
            for k in ("t_cam_obj", "pts", "rays", "depth"):
                g[p + k] = o[k]
            code0 = (np.random.default_rng(100 + i).standard_normal(64) * 0.03).astype(np.float32)
            g[p + "code0"] = code0
            t_oc = torch.inverse(torch.from_numpy(o["t_cam_obj"].copy()))
            zc = torch.from_numpy(code0.copy())
            # loss terms at the initial state (loss.py:22-152)
            jp, jc, rs = ref.loss.compute_sdf_loss(dec, torch.from_numpy(o["pts"]), t_oc, zc)
            g[p + "sdf_Jpose"], g[p + "sdf_Jcode"], g[p + "sdf_res"] = \
                jp[:, 0].numpy(), jc[:, 0].numpy(), rs.reshape(-1).numpy()
            t_co = torch.inverse(t_oc)
            sc = torch.det(t_co[:3, :3]) ** (1 / 3)
            dmin, dmax = t_co[2, 3] - sc, t_co[2, 3] + sc
            sampled = torch.linspace(dmin, dmax, 50)
            K = o["depth"].shape[0]
            dobs = torch.cat([torch.from_numpy(o["depth"]), torch.full((o["rays"].shape[0] - K,), float(1.1 * dmax))])
            jp, jc, rd = ref.loss.compute_render_loss(dec, torch.from_numpy(o["rays"]), dobs, t_oc, sampled, zc, th=0.01)
            g[p + "ren_Jpose"], g[p + "ren_Jcode"], g[p + "ren_res"] = \
                jp[:, 0].numpy(), jc[:, 0].numpy(), rd.reshape(-1).numpy()
            g[p + "sampled"] = sampled.numpy()
            jr, rr_ = ref.loss.compute_rotation_loss_sim3(t_oc)
            g[p + "rot_J"], g[p + "rot_res"] = jr.numpy(), np.float32(rr_)
            # exactly one GN iteration from (t_cam_obj, code0), then the full loop from zero code
            r1 = opt1.reconstruct_object(o["t_cam_obj"].copy(), o["pts"], o["rays"], o["depth"], code0.copy())
            g[p + "it1_t"], g[p + "it1_code"], g[p + "it1_loss"] = r1.t_cam_obj, r1.code, np.float32(r1.loss)
            rN = optN.reconstruct_object(o["t_cam_obj"].copy(), o["pts"], o["rays"], o["depth"])
            g[p + "full_t"], g[p + "full_code"], g[p + "full_loss"] = rN.t_cam_obj, rN.code, np.float32(rN.loss)
            g[p + "full_good"] = np.bool_(rN.is_good)
            if style == "kitti":
                tse = o["t_cam_obj_gt"].copy().astype(np.float32)
                s = np.float32(np.cbrt(np.linalg.det(tse[:3, :3].astype(np.float64))))
                tse[:3, :3] /= s
                g[p + "po_in"], g[p + "po_scale"] = tse.copy(), s
                g[p + "po_out"] = optN.estimate_pose_cam_obj(tse.copy(), float(s), o["pts"], code0.copy()).numpy()
                optN.num_iterations_pose_only = 1
                g[p + "po_out1"] = optN.estimate_pose_cam_obj(tse.copy(), float(s), o["pts"], code0.copy()).numpy()
                optN.num_iterations_pose_only = 5
    # failure case: rays that miss the unit ball entirely -> is_good False (loss.py:73-74)
    o = objs[0]
    far = o["rays"].copy()
    far[:, 0] += 5.0
    rF = optN.reconstruct_object(o["t_cam_obj"].copy(), o["pts"], far, o["depth"])
    g["fail_good"], g["fail_loss"] = np.bool_(rF.is_good), np.float32(rF.loss)
    g["fail_none"] = np.bool_(rF.t_cam_obj is None and rF.code is None)

    np.savez_compressed(os.path.join(OUT, "reference_outputs.npz"), **g)
    print("wrote", os.path.join(OUT, "reference_outputs.npz"), len(g), "arrays")


if __name__ == "__main__":
    main()
