"""The entry points of the reference's ``reconstruct/loss_utils.py``, with the reference's names, argument order and
result shapes, on the native kernels (decoder) and on the device functions of the Gauss-Newton kernels (the small
helpers: sdf_to_occupancy, exp_sim3 / exp_se3, get_points_to_pose_jacobian_sim3 / _se3, get_robust_res,
huber_norm_weights - host arrays in and out through ``qsp_exp_map`` etc. of include/qsp_recon.h):

  decode_sdf(decoder, lat_vec, x, max_batch)        loss_utils.py:51-79   -> (N,)  SDF
  get_batch_sdf_jacobian(decoder, lat_vec, x, 1)    loss_utils.py:82-103  -> (N,1,1) SDF, (N,1,67) d sdf / d [code, xyz]

``decoder`` is a ``DeviceDecoder`` or anything ``as_device_decoder`` accepts (a ``deep_sdf`` Decoder module or its
state dict).  ``lat_vec`` / ``x`` may be numpy arrays or torch tensors; CUDA tensors stay on the device, and the
results come back in the kind of container ``x`` came in.  There is no CPU path.
"""
from __future__ import annotations

import numpy as np

from .. import _native as N
from ..decoder import as_device_decoder


def _is_cuda_tensor(a) -> bool:
    return hasattr(a, "is_cuda") and bool(a.is_cuda)


def _to_numpy(a):
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    return N.f32(a)


def decode_sdf(decoder, lat_vec, x, max_batch: int = 64 ** 3):
    """SDF of the (N,3+) points ``x`` under the code ``lat_vec`` (code_len,) -> (N,).  ``max_batch`` is accepted for
    signature compatibility; the kernel needs no chunking."""
    dev = as_device_decoder(decoder)
    if _is_cuda_tensor(x):
        import torch
        pts = x[:, 0:3].contiguous().float()
        code = torch.as_tensor(lat_vec, dtype=torch.float32, device=x.device).reshape(-1)[:N.CODE_LEN].contiguous()
        out = torch.empty((1, pts.shape[0]), dtype=torch.float32, device=x.device)
        if pts.shape[0]:
            dev.decode_sdf_device(code, pts, out)
        return out[0]
    s = dev.decode_sdf(_to_numpy(lat_vec).reshape(-1)[:N.CODE_LEN], _to_numpy(x)[:, 0:3])
    if hasattr(x, "detach"):
        import torch
        return torch.from_numpy(s)
    return s


def get_batch_sdf_jacobian(decoder, lat_vec, x, out_dim: int = 1):
    """(y (N,1,1), d y / d [code, xyz] (N,1,code_len+3)) as the reference's reverse sweep seeded with ones gives them."""
    if out_dim != 1:
        raise ValueError("the DeepSDF decoder has one output: out_dim must be 1")
    dev = as_device_decoder(decoder)
    if _is_cuda_tensor(x):
        import torch
        pts = x[:, 0:3].contiguous().float()
        n = pts.shape[0]
        code = torch.as_tensor(lat_vec, dtype=torch.float32, device=x.device).reshape(-1)[:N.CODE_LEN].contiguous()
        sdf = torch.empty(n, dtype=torch.float32, device=x.device)
        jac = torch.zeros((n, N.JAC_STRIDE), dtype=torch.float32, device=x.device)
        if n:
            dev.decode_sdf_jacobian_device(code, pts, sdf, jac)
        return sdf.reshape(n, 1, 1), jac[:, :N.CODE_LEN + 3].reshape(n, 1, N.CODE_LEN + 3)
    s, j = dev.decode_sdf_jacobian(_to_numpy(lat_vec).reshape(-1)[:N.CODE_LEN], _to_numpy(x)[:, 0:3])
    s, j = s.reshape(-1, 1, 1), np.ascontiguousarray(j).reshape(-1, 1, N.CODE_LEN + 3)
    if hasattr(x, "detach"):
        import torch
        return torch.from_numpy(s), torch.from_numpy(j)
    return s, j


# ---- the small helpers (loss_utils.py:40-48, 107-265): the device functions of the Gauss-Newton kernels on their own ----
def _back(ref, a: np.ndarray):
    """numpy result in the container kind of ``ref`` (torch tensor on its device, or numpy)."""
    if hasattr(ref, "detach"):
        import torch
        return torch.from_numpy(a).to(ref.device)
    return a


def sdf_to_occupancy(sdf_tensor, th: float = 0.015, device: int = 0):
    """o = 0.5 - clamp(sdf, -th, th) / (2 th): o(x > th) = 0, o(x < -th) = 1 (loss_utils.py:40-48)."""
    s = _to_numpy(sdf_tensor)
    out = np.empty_like(s)
    N.check(N.lib().qsp_sdf_to_occupancy(int(device), N.ptr(s), s.size, float(th), N.ptr(out)))
    return _back(sdf_tensor, out)


def exp_sim3(x, device: int = 0):
    """Exponential map of a sim(3) vector (t, omega, sigma) (7,) -> (4,4) (loss_utils.py:188-233)."""
    v = _to_numpy(x).reshape(-1)
    if v.size != 7:
        raise ValueError("exp_sim3 takes 7 entries")
    out = np.empty((4, 4), np.float32)
    N.check(N.lib().qsp_exp_map(int(device), N.ptr(v), 1, 0, N.ptr(out)))
    return _back(x, out)


def exp_se3(x, device: int = 0):
    """Exponential map of an se(3) vector (t, omega) (6,) -> (4,4) (loss_utils.py:129-163)."""
    v = _to_numpy(x).reshape(-1)
    if v.size != 6:
        raise ValueError("exp_se3 takes 6 entries")
    out = np.empty((4, 4), np.float32)
    N.check(N.lib().qsp_exp_map(int(device), N.ptr(v), 1, 1, N.ptr(out)))
    return _back(x, out)


def get_points_to_pose_jacobian_sim3(points, device: int = 0):
    """[I | -[x]x | x] per transformed point: (n,3) -> (n,3,7), columns (t, omega, sigma) (loss_utils.py:166-185)."""
    p = _to_numpy(points).reshape(-1, 3)
    out = np.empty((p.shape[0], 3, 7), np.float32)
    N.check(N.lib().qsp_point_pose_jacobian(int(device), N.ptr(p), p.shape[0], 0, N.ptr(out)))
    return _back(points, out)


def get_points_to_pose_jacobian_se3(points, device: int = 0):
    """[I | -[x]x] per transformed point: (n,3) -> (n,3,6) (loss_utils.py:107-126)."""
    p = _to_numpy(points).reshape(-1, 3)
    out = np.empty((p.shape[0], 3, 6), np.float32)
    N.check(N.lib().qsp_point_pose_jacobian(int(device), N.ptr(p), p.shape[0], 1, N.ptr(out)))
    return _back(points, out)


def get_robust_res(res, b, device: int = 0):
    """Huber re-weighting of the residuals (loss_utils.py:250-265): -> robust_res (n,1,1), loss = mean(robust^2), w (n,1,1)."""
    r = _to_numpy(res).reshape(-1)
    rob, w, loss = np.empty_like(r), np.empty_like(r), np.empty(1, np.float32)
    N.check(N.lib().qsp_robust_res(int(device), N.ptr(r), r.size, float(b), N.ptr(rob), N.ptr(w), N.ptr(loss)))
    lo = _back(res, loss.reshape(())) if hasattr(res, "detach") else float(loss[0])
    return _back(res, rob.reshape(-1, 1, 1)), lo, _back(res, w.reshape(-1, 1, 1))


def huber_norm_weights(x, b: float = 0.02, device: int = 0):
    """Huber weights of residual norms x >= 0 (loss_utils.py:236-247); unlike the reference ``x`` is not modified."""
    r = _to_numpy(x)
    rob, w, loss = np.empty(r.size, np.float32), np.empty(r.size, np.float32), np.empty(1, np.float32)
    N.check(N.lib().qsp_robust_res(int(device), N.ptr(np.ascontiguousarray(r.reshape(-1))), r.size, float(b), N.ptr(rob), N.ptr(w), N.ptr(loss)))
    return _back(x, w.reshape(r.shape))
