"""The three loss terms of the reference's ``reconstruct/loss.py`` under its names, argument order and result shapes,
computed by the native kernels (there is no CPU path):

  compute_sdf_loss(decoder, pts_surface_cam, t_obj_cam, latent_vector)                        loss.py:22-43
      -> J_pose (M,1,7), J_code (M,1,64), res (M,1,1)
  compute_render_loss(decoder, ray_directions, depth_obs, t_obj_cam, sampled_ray_depth, latent_vector, th)
      -> None (fewer than 10 in-ball samples, loss.py:73-74) or J_pose (P,1,7), J_code (P,1,64), res (P,1,1),
         rows in (ray, sample) order                                                           loss.py:46-152
  compute_rotation_loss_sim3(t_obj_cam) -> J (7,), res                                         loss.py:155-178

Inside ``Optimizer.reconstruct_object`` these terms never leave the GPU: the Jacobian rows are produced by the decoder
kernel and consumed by ``k_accumulate``.  The functions here run ONE recorded iteration of that same launch sequence
(``qsp_set_debug`` + ``qsp_reconstruct_batch``) and read the rows back (``qsp_get_loss_rows``), so what they return is
what the loop accumulates - they exist for callers and parity tests that want a term on its own, not for speed.

``t_obj_cam`` is the camera->object Sim(3) matrix (the reference's ``t_obj_cam``, optimizer.py:126).  The native loop
derives the depth samples from the pose (optimizer.py:144-153: ``linspace(d_min, d_max, S)``), so
``compute_render_loss`` accepts only a ``sampled_ray_depth`` that IS that linspace (as in every call of the reference)
and raises ``ValueError`` otherwise.  Inputs may be numpy arrays or torch tensors; torch in -> torch out (same device).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _native as N
from ..decoder import as_device_decoder


def _np(a):
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return N.f32(a)


def _like(ref, *arrs):
    """numpy results in the container kind of ``ref`` (torch tensor on its device, or numpy)."""
    if hasattr(ref, "detach"):
        import torch
        return tuple(torch.from_numpy(a).to(ref.device) if isinstance(a, np.ndarray) else a for a in arrs)
    return arrs


def _recorded_iteration(decoder, t_obj_cam, pts, rays, depth, code, samples: int, th: float):
    """One recorded Gauss-Newton iteration of a single object; returns (handle, counts)."""
    dev = as_device_decoder(decoder)
    t_cam_obj = np.linalg.inv(_np(t_obj_cam).astype(np.float64)).astype(np.float32).reshape(16)   # optimizer.py:125-126 backwards
    pts = _np(pts).reshape(-1, 3)
    rays = _np(rays).reshape(-1, 3)
    depth = _np(depth).reshape(-1)
    z = np.ascontiguousarray(_np(code).reshape(-1)[:N.CODE_LEN][None])
    off = lambda n: np.array([0, n], np.int32)
    if depth.size == 0:
        depth_buf = np.zeros(1, np.float32)
    else:
        depth_buf = depth
    # weights and Huber deltas do not enter the rows (they are applied by k_accumulate / k_solve); one iteration
    prm = N.GnParams(1.0, 1.0, 0.0, 0.0, 1e30, 1e30, 0.0, 0.0, float(th), 1, int(samples), 0)
    rec = np.empty((1, N.RECORD_LEN), np.float32)
    L = N.lib()
    N.check(L.qsp_set_debug(dev.handle, 1))
    N.check(L.qsp_reconstruct_batch(dev.handle, C.byref(prm), 1, N.ptr(t_cam_obj), N.ptr(z), N.ptr(pts),
                                    N.i32p(off(pts.shape[0])), N.ptr(rays), N.i32p(off(rays.shape[0])),
                                    N.ptr(depth_buf), N.i32p(off(depth.shape[0])), N.ptr(rec), N.FLAG_HOST_PTRS, None))
    cnt = np.empty(6, np.int32)
    N.check(L.qsp_get_iteration(dev.handle, 0, 0, None, None, None, N.ptr(cnt)))
    return dev, cnt


def _rows(dev, term: int) -> np.ndarray:
    L = N.lib()
    n = N.check(L.qsp_get_loss_rows(dev.handle, 0, 0, term, None, -1))
    out = np.empty((max(n, 1), 72), np.float32)
    N.check(L.qsp_get_loss_rows(dev.handle, 0, 0, term, N.ptr(out), n))
    return out[:n]


def _split(rows: np.ndarray):
    return (np.ascontiguousarray(rows[:, None, 0:7]), np.ascontiguousarray(rows[:, None, 7:71]),
            np.ascontiguousarray(rows[:, None, 71:72]))


def compute_sdf_loss(decoder, pts_surface_cam, t_obj_cam, latent_vector):
    """
    :param decoder: DeepSDF decoder (module, state dict or DeviceDecoder)
    :param pts_surface_cam: surface points under the camera frame (M, 3)
    :param t_obj_cam: c2o transformation (4, 4) in Sim(3)
    :param latent_vector: shape code (code_len,)
    :return: Jacobian wrt pose (M, 1, 7), Jacobian wrt shape code (M, 1, 64), sdf residual (M, 1, 1)
    """
    e3 = np.zeros((0, 3), np.float32)
    dev, _ = _recorded_iteration(decoder, t_obj_cam, pts_surface_cam, e3, np.zeros(0, np.float32), latent_vector, 2, 0.01)
    try:
        out = _split(_rows(dev, 0))
    finally:
        N.check(N.lib().qsp_set_debug(dev.handle, 0))
    return _like(pts_surface_cam, *out)


def compute_render_loss(decoder, ray_directions, depth_obs, t_obj_cam, sampled_ray_depth, latent_vector, th=0.01):
    """
    :param decoder: DeepSDF decoder
    :param ray_directions: (N, 3) under the camera frame
    :param depth_obs: (N,) observed depth of every ray (background rays already set to 1.1 * d_max, optimizer.py:153)
    :param t_obj_cam: c2o transformation (4, 4) in Sim(3)
    :param sampled_ray_depth: (S,) depths along the rays; must be linspace(d_min, d_max, S) of the pose
    :param latent_vector: shape code (code_len,)
    :param th: cut-off threshold of the occupancy band
    :return: None, or Jacobian wrt pose (P, 1, 7), wrt code (P, 1, 64), clamped depth residual (P, 1, 1)
    """
    T = _np(t_obj_cam).astype(np.float64)
    T_co = np.linalg.inv(T)
    scale = np.cbrt(np.linalg.det(T_co[:3, :3]))
    d_min, d_max = T_co[2, 3] - scale, T_co[2, 3] + scale
    smp = _np(sampled_ray_depth).reshape(-1)
    S = smp.shape[0]
    if S < 2 or np.abs(smp - np.linspace(d_min, d_max, S)).max() > 1e-4 * max(abs(d_min), abs(d_max), 1.0):
        raise ValueError("sampled_ray_depth must be linspace(d_min, d_max, S) of the pose (optimizer.py:144-151): the "
                         "native loop derives the samples from the pose")
    dev, cnt = _recorded_iteration(decoder, t_obj_cam, np.zeros((0, 3), np.float32), ray_directions, depth_obs,
                                   latent_vector, S, th)
    try:
        if int(cnt[0]) < 10:                                   # loss.py:73-74
            return None
        out = _split(_rows(dev, 1))
    finally:
        N.check(N.lib().qsp_set_debug(dev.handle, 0))
    return _like(ray_directions, *out)


def compute_rotation_loss_sim3(t_obj_cam, device: int = 0):
    """
    :param t_obj_cam: c2o transformation (4, 4) in Sim(3)
    :return: Jacobian (7,) and residual of the gravity-direction prior; (zeros(7), 0.) when the residual is < 1e-7
    """
    T = np.ascontiguousarray(_np(t_obj_cam).reshape(16))
    out = np.empty(8, np.float32)
    N.check(N.lib().qsp_rotation_prior(int(device), N.ptr(T), N.ptr(out)))
    J, = _like(t_obj_cam, out[:7].copy())
    return J, float(out[7])
