"""Synthetic workload generator for the object-reconstruction hot path.

Everything here is numpy-only and seeded with ``numpy.random.default_rng`` (PCG64, stable
across numpy versions), so the same decoder weights and the same object geometry can be
regenerated bit-for-bit in this container (where the reference runs) and on the GPU box
(where it does not).  Nothing is read from disk or from ``/root/reference``.

What it produces mirrors the input distribution of the reference's KITTI front-end
(``reconstruct/kitti_sequence.py:99-220``): per object an object->camera Sim(3) ``t_cam_obj``,
``surface_points`` (M,3) in the camera frame, ``rays`` (N,3) with unit z-component
(foreground rays through surface points first, then background rays) and ``depth`` (K,)
for the foreground rays.

The decoder state dict uses the reference's key names and shapes
(``deep_sdf/deep_sdf_decoder.py:41-63``): ``lin{0..7}.weight_g`` (out,1), ``lin{0..7}.weight_v``
(out,in), ``lin{0..7}.bias`` (out,), ``lin8.weight`` (1,512), ``lin8.bias`` (1,).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import numpy as np

CODE_LEN = 64
HIDDEN = 512
IN_DIM = CODE_LEN + 3
# (in, out) of the nine linear layers; lin3 emits 512-67 because lin4 re-reads the input.
LAYER_DIMS = [(IN_DIM, HIDDEN), (HIDDEN, HIDDEN), (HIDDEN, HIDDEN), (HIDDEN, HIDDEN - IN_DIM),
              (HIDDEN, HIDDEN), (HIDDEN, HIDDEN), (HIDDEN, HIDDEN), (HIDDEN, HIDDEN), (HIDDEN, 1)]


# --------------------------------------------------------------------------------------
# decoder
# --------------------------------------------------------------------------------------
def effective_weights(state: Dict[str, np.ndarray]) -> List[np.ndarray]:
    """Fold weight norm: W = g * v / ||v||_row  (torch ``weight_norm`` with dim=0)."""
    ws = []
    for l in range(8):
        v = state[f"lin{l}.weight_v"].astype(np.float32)
        g = state[f"lin{l}.weight_g"].astype(np.float32)
        nrm = np.sqrt((v.astype(np.float64) ** 2).sum(1, keepdims=True)).astype(np.float32)
        ws.append((g / nrm) * v)
    ws.append(state["lin8.weight"].astype(np.float32))
    return ws


def decode_np(state: Dict[str, np.ndarray], code: np.ndarray, x: np.ndarray,
              chunk: int = 65536) -> np.ndarray:
    """Plain numpy fp32 forward of the decoder; used only to build synthetic inputs."""
    ws = effective_weights(state)
    bs = [state[f"lin{l}.bias"].astype(np.float32) for l in range(9)]
    out = np.empty(x.shape[0], np.float32)
    code = np.asarray(code, np.float32)
    for h in range(0, x.shape[0], chunk):
        xs = x[h:h + chunk].astype(np.float32)
        inp = np.concatenate([np.broadcast_to(code, (xs.shape[0], CODE_LEN)), xs], 1)
        a = inp
        for l in range(9):
            if l == 4:
                a = np.concatenate([a, inp], 1)
            a = a @ ws[l].T + bs[l]
            if l < 8:
                np.maximum(a, 0, out=a)
        out[h:h + chunk] = np.tanh(a[:, 0])
    return out


def _ball_points(rng: np.random.Generator, n: int, radius: float = 1.0) -> np.ndarray:
    d = rng.standard_normal((n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = radius * rng.random(n) ** (1.0 / 3.0)
    return (d * r[:, None]).astype(np.float32)


def make_decoder_state(seed: int = 0, n_calib: int = 26000) -> Dict[str, np.ndarray]:
    """Random-init, calibrated DeepSDF decoder.

    A default-initialised decoder is degenerate for this path (SDF ~ 0.029 everywhere, no zero
    crossing).  Calibration: torch-style uniform init, weight_g scaled by sqrt(6) (unit gain
    through ReLU), hidden biases zero, lin8 rescaled so the pre-tanh output has std 0.25 over the
    unit ball at the zero code, lin8.bias set so its median is 0.
    """
    rng = np.random.default_rng([seed, 0xD5DF])
    state: Dict[str, np.ndarray] = {}
    for l, (fin, fout) in enumerate(LAYER_DIMS):
        bound = 1.0 / math.sqrt(fin)
        w = rng.uniform(-bound, bound, size=(fout, fin)).astype(np.float32)
        if l < 8:
            g = np.sqrt((w.astype(np.float64) ** 2).sum(1, keepdims=True)).astype(np.float32)
            state[f"lin{l}.weight_g"] = (g * np.float32(math.sqrt(6.0))).astype(np.float32)
            state[f"lin{l}.weight_v"] = w
            state[f"lin{l}.bias"] = np.zeros(fout, np.float32)
        else:
            state["lin8.weight"] = w
            state["lin8.bias"] = np.zeros(1, np.float32)
    # calibrate the head on points in the unit ball at the zero code
    pts = _ball_points(np.random.default_rng([seed, 0xCA11B]), n_calib)
    zero = np.zeros(CODE_LEN, np.float32)
    pre = np.arctanh(np.clip(decode_np(state, zero, pts), -0.999999, 0.999999))
    state["lin8.weight"] = (state["lin8.weight"] * np.float32(0.25 / pre.std())).astype(np.float32)
    pre = np.arctanh(np.clip(decode_np(state, zero, pts), -0.999999, 0.999999))
    state["lin8.bias"] = np.array([-np.median(pre)], np.float32)
    return state


def state_to_torch(state: Dict[str, np.ndarray], prefix: str = ""):
    """The same state dict as torch tensors (optionally with DataParallel's ``module.`` prefix,
    as saved checkpoints carry it: ``deep_sdf/workspace.py:215-219``)."""
    import torch
    return {prefix + k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in state.items()}


# --------------------------------------------------------------------------------------
# object geometry
# --------------------------------------------------------------------------------------
_SURFACE_POOLS: Dict[tuple, np.ndarray] = {}


def surface_pool(state: Dict[str, np.ndarray], key, n: int, code: Optional[np.ndarray] = None,
                 seed: int = 0, decode=None, grid: int = 40, tol: float = 0.003) -> np.ndarray:
    """Points of the decoder's own zero level set: ``|sdf(code, x)| < tol`` and ``||x|| < 0.9``.

    Rejection sampling restricted to the cells of a coarse grid whose corners change sign
    (plain rejection over the ball accepts ~1 % and would dominate set-up time).
    ``decode(code, pts) -> sdf`` may be supplied to run the search on the GPU.
    """
    ck = (key, n, seed, grid, tol)
    if ck in _SURFACE_POOLS:
        return _SURFACE_POOLS[ck]
    if code is None:
        code = np.zeros(CODE_LEN, np.float32)
    if decode is None:
        decode = lambda c, p: decode_np(state, c, p)  # noqa: E731
    rng = np.random.default_rng([seed, 0x5F0C])
    lin = np.linspace(-0.9, 0.9, grid, dtype=np.float32)
    gx, gy, gz = np.meshgrid(lin, lin, lin, indexing="ij")
    gp = np.stack([gx, gy, gz], -1).reshape(-1, 3)
    s = decode(code, gp).reshape(grid, grid, grid)
    sign = s > 0
    cell = np.zeros((grid - 1,) * 3, bool)
    ref = sign[:-1, :-1, :-1]
    for dx in (0, 1):
        for dy in (0, 1):
            for dz in (0, 1):
                cell |= sign[dx:grid - 1 + dx, dy:grid - 1 + dy, dz:grid - 1 + dz] != ref
    cells = np.argwhere(cell)
    if len(cells) == 0:
        raise RuntimeError("decoder has no zero crossing inside the 0.9 ball")
    h = lin[1] - lin[0]
    got: List[np.ndarray] = []
    total = 0
    while total < n:
        pick = cells[rng.integers(0, len(cells), size=max(4 * n, 4096))]
        cand = (lin[0] + (pick + rng.random(pick.shape)) * h).astype(np.float32)
        cand = cand[np.linalg.norm(cand, axis=1) < 0.9]
        sv = decode(code, cand)
        keep = cand[np.abs(sv) < tol]
        got.append(keep)
        total += len(keep)
    pool = np.concatenate(got, 0)[:n].astype(np.float32)
    _SURFACE_POOLS[ck] = pool
    return pool


def rot_y(yaw: float) -> np.ndarray:
    c, s = math.cos(yaw), math.sin(yaw)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], np.float64)


def make_object(pool: np.ndarray, seed: int, n_surface: int, n_fg: int, n_bg: int,
                style: str = "kitti", yaw: Optional[float] = None) -> Dict[str, np.ndarray]:
    """One detection: initial ``t_cam_obj`` (object->camera Sim(3)), ``pts``, ``rays``, ``depth``.

    KITTI style (``kitti_sequence.py:118-121``): R = R_y(yaw) * diag(1,-1,-1) so that object +y
    maps to camera -y (otherwise the k4=1e7 gravity prior sees res_rot=2 and the reference
    diverges), scale ~2, translation ~(1, 0.5, 12).  The initial pose is the ground truth moved by
    (0.15, -0.05, 0.20) m and scaled by 1.05.  Redwood style: scale 0.5-1.0, depth 1.5-3 m.
    """
    rng = np.random.default_rng([seed, 0x0B1EC7])
    if style == "kitti":
        yaw = rng.uniform(-math.pi, math.pi) if yaw is None else yaw
        R = rot_y(yaw) @ np.diag([1.0, -1.0, -1.0])
        scale = 2.0 * rng.uniform(0.9, 1.1)
        t = np.array([1.0, 0.5, 12.0]) + rng.uniform(-1, 1, 3) * np.array([3.0, 0.2, 4.0])
        dt = np.array([0.15, -0.05, 0.20])
    elif style == "redwood":
        yaw = rng.uniform(-math.pi, math.pi) if yaw is None else yaw
        R = rot_y(yaw) @ np.diag([1.0, -1.0, -1.0])
        scale = rng.uniform(0.5, 1.0)
        t = np.array([rng.uniform(-0.3, 0.3), rng.uniform(-0.2, 0.2), rng.uniform(1.5, 3.0)])
        dt = np.array([0.03, -0.01, 0.04])
    else:
        raise ValueError(style)
    n_need = max(n_surface, n_fg)
    idx = rng.choice(len(pool), size=n_need, replace=len(pool) < n_need)
    x_obj = pool[idx].astype(np.float64)
    x_cam = (scale * (R @ x_obj.T)).T + t
    # only the camera-facing half is observed by a depth sensor; keep the generator simple and
    # use all points (the reference front-end does no visibility reasoning either).
    pts = x_cam[:n_surface].astype(np.float32)
    fg = x_cam[:n_fg]
    fg_rays = fg / fg[:, 2:3]
    depth = fg[:, 2].astype(np.float32)
    # background rays: around the projected object, outside its bounding box
    uv = x_cam[:, :2] / x_cam[:, 2:3]
    lo, hi = uv.min(0), uv.max(0)
    c, half = (lo + hi) / 2, (hi - lo) / 2
    bg = np.empty((0, 2))
    while len(bg) < n_bg:
        cand = c + (rng.random((4 * n_bg + 16, 2)) * 2 - 1) * half * 1.6
        out = np.any(np.abs(cand - c) > half * 1.05, axis=1)
        bg = np.concatenate([bg, cand[out]], 0)
    bg = bg[:n_bg]
    bg_rays = np.concatenate([bg, np.ones((n_bg, 1))], 1)
    rays = np.concatenate([fg_rays, bg_rays], 0).astype(np.float32)

    t_gt = np.eye(4)
    t_gt[:3, :3] = scale * R
    t_gt[:3, 3] = t
    t_init = np.eye(4)
    t_init[:3, :3] = 1.05 * scale * R
    t_init[:3, 3] = t + dt
    return dict(t_cam_obj=t_init.astype(np.float32), t_cam_obj_gt=t_gt.astype(np.float32),
                pts=pts, rays=rays, depth=depth)


# --------------------------------------------------------------------------------------
# optimiser configurations (the values of the reference's JSON files)
# --------------------------------------------------------------------------------------
def optimizer_config(style: str = "kitti", num_depth_samples: Optional[int] = None,
                     num_iterations: Optional[int] = None) -> dict:
    """``configs/config_kitti.json:21-40`` and ``configs/config_redwood_chair_01053.json:15-30``."""
    if style == "kitti":
        cfg = dict(data_type="KITTI", voxels_dim=32, optimizer=dict(
            code_len=64, num_depth_samples=50, cut_off_threshold=0.01,
            joint_optim=dict(k1=1.0, k2=100.0, k3=0.25, k4=1e7, b1=0.20, b2=0.025,
                             num_iterations=10, learning_rate=1.0, scale_damping=1.0),
            pose_only_optim=dict(num_iterations=5, learning_rate=1.0)))
    elif style == "redwood":
        cfg = dict(data_type="Redwood", voxels_dim=64, optimizer=dict(
            code_len=64, num_depth_samples=50, cut_off_threshold=0.01,
            joint_optim=dict(k1=10.0, k2=100.0, k3=2.5, k4=0.0, b1=0.20, b2=0.02,
                             num_iterations=5, learning_rate=1.0, scale_damping=100.0)))
    else:
        raise ValueError(style)
    if num_depth_samples is not None:
        cfg["optimizer"]["num_depth_samples"] = num_depth_samples
    if num_iterations is not None:
        cfg["optimizer"]["joint_optim"]["num_iterations"] = num_iterations
    return cfg


# The five workloads of BASELINE.json.
WORKLOADS = {
    "kitti_single": dict(style="kitti", objects=1, n_surface=2000, n_fg=2000, n_bg=200,
                         samples=50, iters=10),
    "kitti_batch32": dict(style="kitti", objects=32, n_surface=2000, n_fg=128, n_bg=128,
                          samples=100, iters=10),
    "redwood_chair": dict(style="redwood", objects=1, n_surface=500, n_fg=3896, n_bg=200,
                          samples=128, iters=10),
    "map1024": dict(style="kitti", objects=1024, n_surface=2000, n_fg=128, n_bg=128,
                    samples=100, iters=10),
}


def make_workload(name: str, state: Dict[str, np.ndarray], decoder_seed: int = 0, objects: Optional[int] = None,
                  first_seed: int = 1000, decode=None, pool_size: int = 16384, **over):
    """Returns ``(config_dict, [object dicts])`` for a named workload."""
    w = dict(WORKLOADS[name])
    w.update(over)
    n_obj = w["objects"] if objects is None else objects
    pool = surface_pool(state, ("dec", decoder_seed), pool_size, seed=decoder_seed, decode=decode)
    objs = [make_object(pool, first_seed + i, w["n_surface"], w["n_fg"], w["n_bg"], w["style"])
            for i in range(n_obj)]
    cfg = optimizer_config(w["style"], num_depth_samples=w["samples"], num_iterations=w["iters"])
    return cfg, objs
