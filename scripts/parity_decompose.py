"""GPU box: where the one-iteration difference between the CUDA path and the oracle comes from.

For every fixture object of tests/golden and both arithmetic modes, one recorded GN iteration gives the CUDA path's
Jacobian rows ([J_pose | J_code | residual], qsp_get_loss_rows) and its H, b, dx.  The rows are pushed through the
oracle's accumulate + solve in fp64 (H = k1/P Jr^T Jr + k2/M Js^T Js + ..., optimizer.py:217-254) and compared with
the fp64 oracle:
    rows   : residual error (mean = bias, rms, max), rows whose gradient differs by > 1e-3 relative (ReLU flips)
    dx_acc : CUDA dx  vs  fp64 solve of the CUDA rows          -> what k_accumulate / k_solve add
    dx_rows: fp64 solve of the CUDA rows  vs  fp64 oracle      -> what the decoder rows add
    dx_noflip: the same with the flipped rows replaced by the oracle's rows
Prints one markdown table."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C

import numpy as np
import torch

from qsp_slam_b200 import synth, _native as N
from qsp_slam_b200.decoder import DeviceDecoder
from qsp_slam_b200.reconstruct.optimizer import Optimizer
from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
from oracle import gn_oracle as G


def rows_of(dev, term):
    L = dev._L
    n = N.check(L.qsp_get_loss_rows(dev.handle, 0, 0, term, None, -1))
    out = np.empty((max(n, 1), 72), np.float32)
    N.check(L.qsp_get_loss_rows(dev.handle, 0, 0, term, N.ptr(out), n))
    return out[:n].astype(np.float64)


def solve64(prm, rows_s, rows_r, code, J_rot, res_rot):
    """optimizer.py:205-254 in fp64 from rows [J (71) | raw residual]."""
    def robust(r, b):
        a = np.abs(r)
        nrm = np.where(a <= b, a * a, 2 * b * a - b * b)
        w = np.sqrt(nrm) / np.where(a == 0, 1.0, a)
        return w * r
    Js, rs = rows_s[:, :71], robust(rows_s[:, 71], prm.b2)
    Jr, rr = rows_r[:, :71], robust(rows_r[:, 71], prm.b1)
    H = prm.k1 * Jr.T @ Jr / Jr.shape[0] + prm.k2 * Js.T @ Js / Js.shape[0]
    b = -prm.k1 * Jr.T @ rr / Jr.shape[0] - prm.k2 * Js.T @ rs / Js.shape[0]
    H[7:, 7:] += prm.k3 * np.eye(64)
    b[7:] -= prm.k3 * code
    H[:7, :7] += prm.k4 * np.outer(J_rot, J_rot)
    b[:7] += prm.k4 * J_rot * res_rot
    H[:7, :7] += np.eye(7)
    H[6, 6] += prm.s_damp
    return H, b, np.linalg.solve(H, b)


def rel(a, b):
    return float(np.abs(a - b).max() / np.abs(b).max())


def main():
    golden = dict(np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "reference_outputs.npz")))
    state = synth.make_decoder_state(0)
    dev = DeviceDecoder(state, 0)
    fd64 = G.FoldedDecoder(state, torch.float64)
    print("| mode | object | M | P | res err mean | res err rms | res err max | flipped rows | dx_acc | dx_rows | dx_noflip | code vs o64 |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|")
    for precision in ("tensor", "fp32"):
        for name, style in (("kitti0", "kitti"), ("kitti1", "kitti"), ("redwood0", "redwood"), ("redwood1", "redwood")):
            p = name + "_"
            cfg = synth.optimizer_config(style, num_depth_samples=50, num_iterations=1)
            prm = G.Params(cfg)
            opt = Optimizer(dev, ForceKeyErrorDict(cfg))
            dev.set_precision(precision)
            opt.set_debug(True)
            out = opt.reconstruct_object(golden[p + "t_cam_obj"], golden[p + "pts"], golden[p + "rays"], golden[p + "depth"], golden[p + "code0"])
            it = opt.get_iteration(0)
            rs_, rr_ = rows_of(dev, 0), rows_of(dev, 1)
            opt.set_debug(False)
            dev.set_precision("tensor")
            d64 = lambda a: torch.from_numpy(np.ascontiguousarray(a)).double()
            t_oc = torch.inverse(d64(golden[p + "t_cam_obj"]))
            o = G.gn_iteration(fd64, prm, t_oc, d64(golden[p + "code0"]), d64(golden[p + "pts"]), d64(golden[p + "rays"]), d64(golden[p + "depth"]))
            os_ = torch.cat([o["sdf"]["J_pose"], o["sdf"]["J_code"], o["sdf"]["res"][:, None]], 1).numpy()
            or_ = torch.cat([o["render"]["J_pose"], o["render"]["J_code"], o["render"]["res"][:, None]], 1).numpy()
            if rr_.shape != or_.shape:
                print(f"| {precision} | {name} | kept sets differ: {rr_.shape[0]} vs {or_.shape[0]} |")
                continue
            eres = rs_[:, 71] - os_[:, 71]
            def flipped(a, b):
                return np.linalg.norm(a[:, :71] - b[:, :71], axis=1) / np.maximum(np.linalg.norm(b[:, :71], axis=1), 1e-30) > 1e-3
            fs, fr = flipped(rs_, os_), flipped(rr_, or_)
            J_rot, res_rot = o["J_rot"].numpy(), float(o["res_rot"])
            code = golden[p + "code0"].astype(np.float64)
            _, _, dx_rows = solve64(prm, rs_, rr_, code, J_rot, res_rot)
            rs2, rr2 = rs_.copy(), rr_.copy()
            rs2[fs], rr2[fr] = os_[fs], or_[fr]
            _, _, dx_nf = solve64(prm, rs2, rr2, code, J_rot, res_rot)
            dx64 = o["dx"].numpy()
            print(f"| {precision} | {name} | {rs_.shape[0]} | {rr_.shape[0]} | {eres.mean():.1e} | {np.sqrt((eres ** 2).mean()):.1e} | "
                  f"{np.abs(eres).max():.1e} | {int(fs.sum())} + {int(fr.sum())} | {rel(it['dx'].astype(np.float64), dx_rows):.1e} | "
                  f"{rel(dx_rows, dx64):.1e} | {rel(dx_nf, dx64):.1e} | "
                  f"{float(np.abs(out.code - o['code'].numpy()).max() / max(np.abs(o['code'].numpy()).max(), 0.05)):.1e} |", flush=True)


if __name__ == "__main__":
    main()
