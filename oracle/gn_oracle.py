"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's object-reconstruction hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import this module, and only as the checker or the timed CPU baseline.  The product
package (``qsp_slam_b200``) never imports it and has no CPU fallback.

Parity status: the reference ships no golden vectors or tests for this path,
so this restatement is pinned against OUTPUTS OF THE REFERENCE ITSELF, produced in the build
container by ``oracle/make_golden.py`` (which imports the unmodified reference through
``oracle/ref_harness.py``) and committed under ``tests/golden/``.  ``tests/test_oracle_golden.py``
checks every function below against those fixtures.

The restatement is explicit torch-CPU tensor arithmetic (fp32 by default, fp64 on request): no
autograd, no ``nn.Module``, weight norm folded once.  Each function cites the reference lines it
follows (paths relative to /root/reference).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import numpy as np
import torch

CODE_LEN = 64


# ----------------------------------------------------------------------------------------------
# decoder  (deep_sdf/deep_sdf_decoder.py:75-110, parameters :41-63)
# ----------------------------------------------------------------------------------------------
class FoldedDecoder:
    """Effective weights W = g * v / ||v||_row (torch weight_norm, dim=0) and biases."""

    def __init__(self, state: Dict[str, np.ndarray], dtype=torch.float32):
        self.dtype = dtype
        self.W: List[torch.Tensor] = []
        self.b: List[torch.Tensor] = []
        for l in range(8):
            v = torch.from_numpy(np.asarray(state[f"lin{l}.weight_v"], np.float32))
            g = torch.from_numpy(np.asarray(state[f"lin{l}.weight_g"], np.float32))
            # torch._weight_norm: v * (g / norm(v)) with the norm over dim 1, in fp32
            w = v * (g / v.norm(2, dim=1, keepdim=True))
            self.W.append(w.to(dtype))
            self.b.append(torch.from_numpy(np.asarray(state[f"lin{l}.bias"], np.float32)).to(dtype))
        self.W.append(torch.from_numpy(np.asarray(state["lin8.weight"], np.float32)).to(dtype))
        self.b.append(torch.from_numpy(np.asarray(state["lin8.bias"], np.float32)).to(dtype))


def decoder_forward(dec: FoldedDecoder, code: torch.Tensor, x: torch.Tensor, keep=None) -> torch.Tensor:
    """SDF of points ``x`` (n,3) under ``code`` (64,).  Input row is [code, xyz]
    (loss_utils.py:71-72); the input is re-concatenated in front of lin4 as [h3, code, xyz]
    (deep_sdf_decoder.py:87-88); ReLU after lin0..lin7, tanh on the output (:103,:107-108)."""
    n = x.shape[0]
    inp = torch.cat([code.to(dec.dtype).expand(n, -1), x.to(dec.dtype)], 1)
    a = inp
    for l in range(9):
        if l == 4:
            a = torch.cat([a, inp], 1)
        a = torch.addmm(dec.b[l], a, dec.W[l].t())
        if l < 8:
            if keep is not None:
                keep.append(a > 0)
            a = torch.relu(a)
    return torch.tanh(a[:, 0])


def decoder_forward_jacobian(dec: FoldedDecoder, code: torch.Tensor, x: torch.Tensor):
    """(sdf (n,), d sdf / d [code, xyz] (n,67)) - what ``get_batch_sdf_jacobian`` returns
    (loss_utils.py:82-103: one reverse sweep seeded with ones), written out by hand."""
    masks: List[torch.Tensor] = []
    s = decoder_forward(dec, code, x, keep=masks)
    g = ((1.0 - s * s)[:, None] * dec.W[8]) * masks[7]          # d/d(pre-activation of lin7)
    skip = None
    for l in range(7, 0, -1):
        g = g @ dec.W[l]                                         # d/d(input of lin l)
        if l == 4:
            skip = g[:, -(CODE_LEN + 3):]
            g = g[:, :-(CODE_LEN + 3)]
        g = g * masks[l - 1]
    g = g @ dec.W[0] + skip
    return s, g


def decode_sdf(dec: FoldedDecoder, code, x, max_batch: int = 64 ** 3) -> torch.Tensor:
    """Chunked no-grad decode (loss_utils.py:51-79)."""
    code = torch.as_tensor(code)
    x = torch.as_tensor(x)
    out = [decoder_forward(dec, code, x[h:h + max_batch]) for h in range(0, x.shape[0], max_batch)]
    return torch.cat(out, 0) if out else torch.zeros(0, dtype=dec.dtype)


def create_voxel_grid(vol_dim: int = 128) -> torch.Tensor:
    """reconstruct/utils.py:98-117, including its true-division quirk: ``index / vol_dim`` is not
    floored under torch >= 1.7, so the x and y columns carry fractional offsets."""
    voxel_size = 2.0 / (vol_dim - 1)
    idx = torch.arange(0, vol_dim ** 3, 1, dtype=torch.int64)
    v = torch.zeros(vol_dim ** 3, 3)
    v[:, 2] = idx % vol_dim
    v[:, 1] = (idx / vol_dim) % vol_dim
    v[:, 0] = ((idx / vol_dim) / vol_dim) % vol_dim
    v[:, 0] = v[:, 0] * voxel_size + (-1)
    v[:, 1] = v[:, 1] * voxel_size + (-1)
    v[:, 2] = v[:, 2] * voxel_size + (-1)
    return v


# ----------------------------------------------------------------------------------------------
# small helpers  (reconstruct/loss_utils.py)
# ----------------------------------------------------------------------------------------------
def points_to_pose_jacobian_sim3(p: torch.Tensor) -> torch.Tensor:
    """(n,3,7) = [ I3 | -[p]x | p ], columns ordered (t, omega, sigma)  (loss_utils.py:166-185).
    The reference stacks the rows of the hat matrix on dim=-1, so they become COLUMNS:
    -[p]x = [[0, z, -y], [-z, 0, x], [y, -x, 0]]."""
    n = p.shape[0]
    J = torch.zeros(n, 3, 7, dtype=p.dtype)
    J[:, 0, 0] = J[:, 1, 1] = J[:, 2, 2] = 1
    x, y, z = p[:, 0], p[:, 1], p[:, 2]
    J[:, 0, 4], J[:, 0, 5] = z, -y
    J[:, 1, 3], J[:, 1, 5] = -z, x
    J[:, 2, 3], J[:, 2, 4] = y, -x
    J[:, :, 6] = p
    return J


def sdf_to_occupancy(s: torch.Tensor, th: float) -> torch.Tensor:
    """loss_utils.py:40-48."""
    return 0.5 - torch.clamp(s, min=-th, max=th) / (2 * th)


def robust_res(res: torch.Tensor, b: float):
    """Huber re-weighting (loss_utils.py:236-265): w = 1 for |r|<=b else sqrt(2b|r|-b^2)/|r|;
    a zero residual gets weight sqrt(0)/1 = 0.  Returns (w*r, mean((w*r)^2), w)."""
    r = res.reshape(-1)
    a = r.abs()
    nrm = torch.where(a <= b, a * a, 2 * b * a - b * b)
    den = torch.where(a == 0, torch.ones_like(a), a)
    w = torch.sqrt(nrm) / den
    rr = w * r
    return rr, (rr * rr).mean(), w


def exp_sim3(x: torch.Tensor) -> torch.Tensor:
    """loss_utils.py:188-233, branch for branch (including c = 0 for sigma <= 1e-8, :223)."""
    dt = x.dtype
    v, w, s = x[:3], x[3:6], x[6]
    wh = torch.zeros(3, 3, dtype=dt)
    wh[0, 1], wh[0, 2], wh[1, 0], wh[1, 2], wh[2, 0], wh[2, 1] = -w[2], w[1], w[2], -w[0], -w[1], w[0]
    wh2 = wh @ wh
    th = torch.norm(w)
    th2 = th ** 2
    sn, cs = torch.sin(th), torch.cos(th)
    es = torch.exp(s)
    s2 = s ** 2
    eye = torch.eye(3, dtype=dt)
    if th <= 1e-8:
        ew = eye
        j = eye if s == 0 else ((es - 1.0) / s) * eye
    else:
        ew = eye + wh * sn / th + wh2 * (1.0 - cs) / th2
        a, b = es * sn, es * cs
        c = 0.0 if s <= 1e-8 else (es - 1.0) / s
        k1 = (a * s + (1 - b) * th) / (s2 + th2)
        k2 = c - ((b - 1) * s + a * th) / (s2 + th2)
        j = c * eye + k1 * wh / th + k2 * wh2 / th2
    out = torch.eye(4, dtype=dt)
    out[:3, :3] = es * ew
    out[:3, 3] = j @ v
    return out


def exp_se3(x: torch.Tensor) -> torch.Tensor:
    """loss_utils.py:129-163."""
    dt = x.dtype
    v, w = x[:3], x[3:6]
    wh = torch.zeros(3, 3, dtype=dt)
    wh[0, 1], wh[0, 2], wh[1, 0], wh[1, 2], wh[2, 0], wh[2, 1] = -w[2], w[1], w[2], -w[0], -w[1], w[0]
    wh2 = wh @ wh
    th = torch.norm(w)
    eye = torch.eye(3, dtype=dt)
    if th <= 1e-8:
        ew, j = eye, eye
    else:
        ew = eye + wh * torch.sin(th) / th + wh2 * (1.0 - torch.cos(th)) / th ** 2
        j = eye + (1 - torch.cos(th)) / th ** 2 * wh + (th - torch.sin(th)) / th ** 3 * wh2
    out = torch.eye(4, dtype=dt)
    out[:3, :3] = ew
    out[:3, 3] = j @ v
    return out


# ----------------------------------------------------------------------------------------------
# loss terms  (reconstruct/loss.py)
# ----------------------------------------------------------------------------------------------
def transform_points(t_obj_cam: torch.Tensor, p_cam: torch.Tensor) -> torch.Tensor:
    """x_o = R x_c + t written as the reference writes it: (x[...,None,:] * R).sum(-1) + t
    (loss.py:31-32, :62-63)."""
    return (p_cam[..., None, :] * t_obj_cam[:3, :3]).sum(-1) + t_obj_cam[:3, 3]


def sdf_term(dec: FoldedDecoder, pts_cam, t_obj_cam, code):
    """compute_sdf_loss (loss.py:22-43): residual is the SDF; J_pose = ds/dx * dx/dT (M,7),
    J_code = ds/dz (M,64)."""
    x_o = transform_points(t_obj_cam, pts_cam)
    s, g = decoder_forward_jacobian(dec, code, x_o)
    J_pose = torch.bmm(g[:, None, CODE_LEN:], points_to_pose_jacobian_sim3(x_o))[:, 0]
    return dict(J_pose=J_pose, J_code=g[:, :CODE_LEN], res=s, x_obj=x_o, grad=g)


def render_term(dec: FoldedDecoder, rays, depth_obs, t_obj_cam, sampled, code, th: float):
    """compute_render_loss (loss.py:46-152).  Returns None when fewer than 10 samples fall inside
    the unit ball (:73-74); otherwise the Jacobians/residuals of the kept band points plus every
    intermediate index set (ray-major order, as torch.where gives it)."""
    dt = dec.dtype
    S = sampled.shape[0]
    x_c = rays[..., None, :] * sampled[:, None]                       # (N,S,3)  :60
    x_o = transform_points(t_obj_cam, x_c)                            # :62-63
    vi, vj = torch.where(torch.norm(x_o, dim=-1) < 1.0)               # :68
    if vi.shape[0] < 10:
        return None
    q = x_o[vi, vj]
    s = decode_sdf(dec, code, q)                                      # :77-78
    occ = torch.zeros(x_o.shape[0], S, dtype=dt)
    occ[vi, vj] = sdf_to_occupancy(s, th)                             # :84-86
    band = (s > -th) & (s < th)                                       # :88
    bi, bj = vi[band], vj[band]
    o_rows = occ[bi]                                                  # (m,S) :93
    m = o_rows.shape[0]
    d_min, d_max = sampled[0], sampled[-1]
    T = torch.cumprod(1 - o_rows, dim=-1)                             # :99
    T_aug = torch.cat([torch.ones(m, 1, dtype=dt), T], -1)
    o_aug = torch.cat([o_rows, torch.ones(m, 1, dtype=dt)], -1)
    d_aug = torch.cat([sampled, (1.1 * d_max).reshape(1)], -1)
    d_u = (d_aug * (o_aug * T_aug)).sum(-1)                           # :112-114
    o_k = occ[bi, bj]
    l = torch.arange(S)[None, :]
    Tm = torch.where(l < bj[:, None], torch.zeros((), dtype=dt), T)   # :121
    de_do = Tm.sum(-1) / (1.0 - o_k)                                  # :122
    nz = de_do > 1e-2                                                 # :125
    delta_d = (d_max - d_min) / (S - 1)
    de_ds = de_do[nz] * delta_d * (-1.0 / (2 * th))                   # :128-130
    ki, kj = bi[nz], bj[nz]
    res = torch.clamp(depth_obs[ki] - d_u[nz], -0.30, 0.30)           # :135-140
    p = x_o[ki, kj]
    if p.shape[0] > 0:
        _, g = decoder_forward_jacobian(dec, code, p)                 # :144
    else:
        g = torch.zeros(0, CODE_LEN + 3, dtype=dt)
    de_di = de_ds[:, None] * g                                        # :145
    J_pose = torch.bmm(de_di[:, None, CODE_LEN:], points_to_pose_jacobian_sim3(p))[:, 0]
    return dict(J_pose=J_pose, J_code=de_di[:, :CODE_LEN], res=res,
                valid_ray=vi, valid_sample=vj, sdf=s, band_ray=bi, band_sample=bj,
                de_do=de_do, d_u=d_u, keep_ray=ki, keep_sample=kj, de_ds=de_ds, x_obj=p, grad=g)


def rotation_term(t_obj_cam: torch.Tensor):
    """compute_rotation_loss_sim3 (loss.py:155-178): res = 1 - (R_co e_y).n_g with e_y=(0,1,0),
    n_g=(0,-1,0); below 1e-7 it returns (zeros(7), 0)."""
    dt = t_obj_cam.dtype
    t_cam_obj = torch.inverse(t_obj_cam)
    r_co = t_cam_obj[:3, :3]
    scale = torch.det(r_co) ** (1 / 3)
    r_co = r_co / scale
    r_oc = torch.inverse(r_co)
    ey = torch.tensor([0.0, 1.0, 0.0], dtype=dt)
    ng = torch.tensor([0.0, -1.0, 0.0], dtype=dt)
    res = 1.0 - torch.dot(r_co @ ey, ng)
    if res < 1e-7:
        return torch.zeros(7, dtype=dt), torch.zeros((), dtype=dt)
    J = torch.zeros(7, dtype=dt)
    J[3:6] = torch.linalg.cross(r_oc @ ng, ey)
    return J, res


# ----------------------------------------------------------------------------------------------
# Gauss-Newton loop  (reconstruct/optimizer.py)
# ----------------------------------------------------------------------------------------------
class Params:
    """The scalars ``Optimizer.__init__`` copies out of the config (optimizer.py:27-44)."""

    def __init__(self, cfg: dict):
        o = cfg["optimizer"]
        j = o["joint_optim"]
        self.k1, self.k2, self.k3, self.k4 = j["k1"], j["k2"], j["k3"], j["k4"]
        self.b1, self.b2 = j["b1"], j["b2"]
        self.lr, self.s_damp = j["learning_rate"], j["scale_damping"]
        self.iters = j["num_iterations"]
        self.code_len = o["code_len"]
        self.S = o["num_depth_samples"]
        self.th = o["cut_off_threshold"]
        self.pose_only_iters = o.get("pose_only_optim", {}).get("num_iterations", 0)


def depth_range(t_obj_cam: torch.Tensor, S: int):
    """optimizer.py:144-151: T_co = inv(T_oc); scale = det(R)^(1/3); d = T_co[2,3] -/+ scale;
    linspace(d_min, d_max, S)."""
    t_cam_obj = torch.inverse(t_obj_cam)
    scale = torch.det(t_cam_obj[:3, :3]) ** (1 / 3)
    d_min, d_max = t_cam_obj[2, 3] - 1.0 * scale, t_cam_obj[2, 3] + 1.0 * scale
    return d_min, d_max, torch.linspace(float(d_min), float(d_max), S, dtype=t_obj_cam.dtype)


def gn_iteration(dec: FoldedDecoder, prm: Params, t_obj_cam, code, pts, rays, depth_fg):
    """One pass of the loop body (optimizer.py:141-263) from a given state.  Returns a dict with
    ``ok`` False at the reference's early exits, else H, b, dx, the new state and the loss."""
    dt = dec.dtype
    K = depth_fg.shape[0]
    d_min, d_max, sampled = depth_range(t_obj_cam, prm.S)
    depth_obs = torch.cat([depth_fg.to(dt), torch.full((rays.shape[0] - K,), float(1.1 * d_max), dtype=dt)])
    st = sdf_term(dec, pts, t_obj_cam, code)
    rs, sdf_loss, _ = robust_res(st["res"], prm.b2)
    if math.isnan(float(sdf_loss)):
        return dict(ok=False, why="sdf_nan")
    rt = render_term(dec, rays, depth_obs, t_obj_cam, sampled, code, prm.th)
    if rt is None:
        return dict(ok=False, why="few_samples")
    rr, render_loss, _ = robust_res(rt["res"], prm.b1)
    if math.isnan(float(render_loss)):
        return dict(ok=False, why="render_nan", render=rt)
    J_rot, res_rot = rotation_term(t_obj_cam)
    loss = prm.k1 * render_loss + prm.k2 * sdf_loss                                  # :205
    Js = torch.cat([st["J_pose"], st["J_code"]], 1)
    Jr = torch.cat([rt["J_pose"], rt["J_code"]], 1)
    H = prm.k1 * (Jr.t() @ Jr) / Jr.shape[0] + prm.k2 * (Js.t() @ Js) / Js.shape[0]  # :218,225,231
    b = -prm.k1 * (Jr.t() @ rr) / Jr.shape[0] - prm.k2 * (Js.t() @ rs) / Js.shape[0]  # :219,226,233
    H[7:, 7:] += prm.k3 * torch.eye(CODE_LEN, dtype=dt)                              # :232
    b[7:] -= prm.k3 * code                                                           # :234
    H[:7, :7] += prm.k4 * torch.outer(J_rot, J_rot)                                  # :241,243
    b[:7] -= prm.k4 * (-(J_rot * res_rot))                                           # :242,244
    H[:7, :7] += torch.eye(7, dtype=dt)                                              # :251
    H[6, 6] += prm.s_damp                                                            # :252
    dx = torch.inverse(H) @ b                                                        # :254
    new_t = exp_sim3(prm.lr * dx[:7]) @ t_obj_cam                                    # :261-262
    new_code = code + prm.lr * dx[7:]                                                # :263
    return dict(ok=True, H=H, b=b, dx=dx, t_obj_cam=new_t, code=new_code, loss=loss,
                sdf_loss=sdf_loss, render_loss=render_loss, sdf=st, render=rt,
                J_rot=J_rot, res_rot=res_rot, d_min=d_min, d_max=d_max)


def reconstruct_object(dec: FoldedDecoder, prm: Params, t_cam_obj, pts, rays, depth, code=None,
                       trace: Optional[list] = None):
    """Optimizer.reconstruct_object (optimizer.py:96-281).  Returns a dict with the reference's
    keys: t_cam_obj (4,4) or None, code (64,) or None, is_good, loss."""
    dt = dec.dtype
    z = torch.zeros(prm.code_len, dtype=dt) if code is None else \
        torch.as_tensor(np.asarray(code)[:prm.code_len].copy()).to(dt)
    t_obj_cam = torch.inverse(torch.as_tensor(np.asarray(t_cam_obj)).to(dt))
    pts, rays, depth = (torch.as_tensor(np.ascontiguousarray(a)).to(dt) for a in (pts, rays, depth))
    loss = 0.0
    for _ in range(prm.iters):
        it = gn_iteration(dec, prm, t_obj_cam, z, pts, rays, depth)
        if not it["ok"]:
            return dict(t_cam_obj=None, code=None, is_good=False, loss=loss)
        if trace is not None:
            trace.append(it)
        t_obj_cam, z, loss = it["t_obj_cam"], it["code"], float(it["loss"])
    return dict(t_cam_obj=torch.inverse(t_obj_cam).numpy(), code=z.numpy(), is_good=True, loss=loss)


def estimate_pose_cam_obj(dec: FoldedDecoder, prm: Params, t_co_se3, scale, pts, code,
                          trace: Optional[list] = None):
    """Optimizer.estimate_pose_cam_obj (optimizer.py:47-93): 6-dof GN on the SDF term with the code
    fixed; H = J^T J / M + 1e-2 I, b uses the RAW residual (:75), Huber(0.05) only feeds the unused
    loss (:70); inliers |res| <= 0.05 are kept after iteration index 4 (:80-82)."""
    dt = dec.dtype
    t = torch.as_tensor(np.asarray(t_co_se3).copy()).to(dt)
    t[:3, :3] *= scale
    t_obj_cam = torch.inverse(t)
    z = torch.as_tensor(np.asarray(code).copy()).to(dt)
    p = torch.as_tensor(np.ascontiguousarray(pts)).to(dt)
    for e in range(prm.pose_only_iters):
        st = sdf_term(dec, p, t_obj_cam, z)
        J = st["J_pose"][:, :6]
        H = (J.t() @ J) / J.shape[0] + 1e-2 * torch.eye(6, dtype=dt)
        b = -(J.t() @ st["res"]) / J.shape[0]
        dx = torch.inverse(H) @ b
        t_obj_cam = exp_se3(dx) @ t_obj_cam
        if trace is not None:
            trace.append(dict(H=H, b=b, dx=dx, t_obj_cam=t_obj_cam, n=J.shape[0]))
        if e == 4:
            p = p[st["res"].abs() <= 0.05]
    out = torch.inverse(t_obj_cam)
    out[:3, :3] /= scale
    return out.numpy()
