"""The two decoder entry points of the reference's ``reconstruct/loss_utils.py`` that callers outside the
optimiser use, with the reference's names, argument order and result shapes, on the native kernels:

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
