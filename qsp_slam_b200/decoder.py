"""Device-resident DeepSDF decoder handle.

Takes the UNCHANGED deep_sdf state dict (``lin{l}.weight_g / weight_v / bias``, ``lin8.weight/bias``;
deep_sdf/deep_sdf_decoder.py:41-63), with or without DataParallel's ``module.`` prefix
(deep_sdf/workspace.py:215-219), folds weight norm inside the native library and keeps the operands
on one GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import numpy as np

from . import _native as N

_IN = [67, 512, 512, 512, 512, 512, 512, 512, 512]
_OUT = [512, 512, 512, 445, 512, 512, 512, 512, 1]


def _to_numpy_state(obj) -> Dict[str, np.ndarray]:
    if hasattr(obj, "state_dict"):
        obj = obj.state_dict()
    out = {}
    for k, v in dict(obj).items():
        if k.startswith("module."):
            k = k[len("module."):]
        if hasattr(v, "detach"):
            v = v.detach().cpu().numpy()
        out[k] = np.ascontiguousarray(v, dtype=np.float32)
    return out


class DeviceDecoder:
    """Owns a ``qsp_decoder*``.  ``decoder`` may be a state dict, an ``nn.Module`` or another
    ``DeviceDecoder`` (returned as is by :func:`as_device_decoder`)."""

    def __init__(self, decoder, device: int = 0, calibrate: Optional[bool] = None):
        """``calibrate`` (default: on, ``QSP_CALIBRATE=0`` turns it off): measure the fp16 first-pass error on THESE
        weights and raise the candidate margin if the defaults - measured on random-init decoders - do not cover it
        (:meth:`calibrate_first_pass_margin`; ~50 ms, a few decoder launches)."""
        st = _to_numpy_state(decoder)
        for l in range(8):
            for key, shape in ((f"lin{l}.weight_v", (_OUT[l], _IN[l])), (f"lin{l}.weight_g", (_OUT[l], 1)),
                               (f"lin{l}.bias", (_OUT[l],))):
                if key not in st:
                    raise KeyError(f"decoder state dict lacks '{key}' (expected the 8x512 weight-normed DeepSDF layout)")
                if st[key].size != int(np.prod(shape)):
                    raise ValueError(f"'{key}' has shape {st[key].shape}, expected {shape}")
        if st["lin8.weight"].size != 512 or st["lin8.bias"].size != 1:
            raise ValueError("lin8 must be 512 -> 1")
        w = N.DecoderWeights()
        self._keep = st
        fp = C.POINTER(C.c_float)
        for l in range(8):
            w.weight_v[l] = st[f"lin{l}.weight_v"].ctypes.data_as(fp)
            w.weight_g[l] = st[f"lin{l}.weight_g"].ctypes.data_as(fp)
            w.bias[l] = st[f"lin{l}.bias"].ctypes.data_as(fp)
        w.lin8_weight = st["lin8.weight"].ctypes.data_as(fp)
        w.lin8_bias = st["lin8.bias"].ctypes.data_as(fp)
        h = C.c_void_p()
        self._L = N.lib()                     # the library build this handle belongs to (tests may bind another one)
        N.check(self._L.qsp_decoder_create(C.byref(w), int(device), C.byref(h)))
        self.handle = h
        self.device = int(device)
        self.state = st
        self.margin_calibration = None
        if calibrate is None:
            calibrate = os.environ.get("QSP_CALIBRATE", "1") != "0"
        if calibrate:
            self.margin_calibration = self.calibrate_first_pass_margin(n_points=131072, codes_per_norm=2)

    def close(self):
        if getattr(self, "handle", None):
            self._L.qsp_decoder_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # nn.Module look-alikes so that reference-style call sites keep working
    def eval(self):
        return self

    def cuda(self, *a, **k):
        return self

    # --- kernel (a) -------------------------------------------------------------------------------
    def set_precision(self, precision: str):
        """Arithmetic of the decoder inside the GN loop: "tensor" (tcgen05, fp16-pair operands, default)
        or "fp32" (exact-fp32 CUDA-core kernel, for validation)."""
        N.check(self._L.qsp_set_precision(self.handle, {"tensor": N.PREC_TENSOR, "fp32": N.PREC_FP32_CORES}[precision]))

    def set_first_pass_margin(self, margin0: float = 0.0045, margin_per_norm: float = 0.004):
        """Samples whose fp16 first-pass |sdf| is below cut_off + margin0 + margin_per_norm * |code| are re-evaluated
        with fp32-equivalent arithmetic (GN loop and mesh extraction); results are exact while the first-pass error
        stays below the margin."""
        N.check(self._L.qsp_set_first_pass_margin(self.handle, float(margin0), float(margin_per_norm)))

    def calibrate_first_pass_margin(self, code_norms=(0.0, 0.4, 0.8, 1.6, 2.4), n_points: int = 262144, codes_per_norm: int = 4,
                                    safety: float = 1.5, seed: int = 0, allow_lower: bool = False) -> dict:
        """Measures the error of the fp16 first pass against the fp32-equivalent kernel on THIS decoder's weights and
        sets the margin from it (a trained decoder may have larger activations than the random-init decoders the
        defaults were measured on, ``scripts/fast_error_tail.py``).  For every code norm in ``code_norms``,
        ``codes_per_norm`` random codes of that norm are decoded at ``n_points`` uniform points of the unit ball in
        both modes; margin0 = safety * max error at norm 0, margin_per_norm = safety * the largest slope
        (err(norm) - err(0)) / norm.  The margins are never set below the defaults unless ``allow_lower``.  Returns
        the measured maxima and the margins in force.  Runs entirely on the device (two decoder launches per norm)."""
        rng = np.random.default_rng(seed)
        x = rng.standard_normal((n_points, 3))
        x = (x / np.linalg.norm(x, axis=1, keepdims=True) * rng.random((n_points, 1)) ** (1.0 / 3.0)).astype(np.float32)
        err = {}
        for nz in code_norms:
            z = rng.standard_normal((codes_per_norm, N.CODE_LEN))
            z = (z / np.linalg.norm(z, axis=1, keepdims=True) * nz).astype(np.float32)
            if nz == 0.0:
                z = z[:1]
            fast = self.decode_sdf(z, x, mode="fast")
            exact = self.decode_sdf(z, x, mode="tensor")
            err[float(nz)] = float(np.abs(fast - exact).max())
        e0 = err.get(0.0, min(err.values()))
        slope = max([0.0] + [(e - e0) / nz for nz, e in err.items() if nz > 0.0])
        m0, m1 = safety * e0, safety * slope
        if not allow_lower:
            m0, m1 = max(m0, 0.0045), max(m1, 0.004)
        self.set_first_pass_margin(m0, m1)
        return dict(max_error_by_code_norm=err, margin0=m0, margin_per_norm=m1)

    def call_stats(self) -> dict:
        """Diagnostics of the last GN call on this handle (``qsp_get_call_stats``): ``overflow`` (objects failed for
        lack of candidate-row capacity), ``max_first_pass_error`` and ``n_margin_excursions`` (re-evaluated samples
        whose fp16 first-pass error exceeded 3/4 of the margin)."""
        ov, ex, er = C.c_int32(), C.c_int32(), C.c_float()
        N.check(self._L.qsp_get_call_stats(self.handle, C.byref(ov), C.byref(er), C.byref(ex)))
        return dict(overflow=int(ov.value), max_first_pass_error=float(er.value), n_margin_excursions=int(ex.value))

    def check_guards(self) -> int:
        """Debug (QSP_GUARD=1 in the environment): number of workspace buffers whose canary bands were overwritten
        since they were allocated, -1 if none is guarded; raises with the buffer names if any was."""
        n = self._L.qsp_check_guards(self.handle)
        if n > 0:
            raise N.NativeError(self._L.qsp_last_error().decode())
        return n

    def decode_sdf(self, codes, pts, mode: str = "tensor") -> np.ndarray:
        """codes (C,64) or (64,), pts (n,3) host arrays -> (C,n) or (n,) float32  (loss_utils.py:51-79).
        mode: "tensor" (fp16-pair tensor-core kernel), "fast" (single-pass fp16), "fp32" (CUDA cores)."""
        codes = N.f32(codes)
        single = codes.ndim == 1
        codes = codes.reshape(-1, 64)
        pts = N.f32(pts)[:, :3]
        pts = np.ascontiguousarray(pts)
        out = np.empty((codes.shape[0], pts.shape[0]), np.float32)
        if pts.shape[0]:
            N.check(self._L.qsp_decode_sdf(self.handle, N.ptr(codes), codes.shape[0], N.ptr(pts), pts.shape[0],
                                           N.ptr(out), N.FLAG_HOST_PTRS | N.MODE_FLAGS[mode], None))
        return out[0] if single else out

    def decode_sdf_device(self, codes, pts, out, sync: bool = True, mode: str = "tensor"):
        """Same with torch CUDA tensors (float32, contiguous); ``out`` (C,n) is written in place."""
        n_codes = codes.numel() // 64
        n = pts.shape[0]
        N.check(self._L.qsp_decode_sdf(self.handle, N.ptr(codes), n_codes, N.ptr(pts), n, N.ptr(out),
                                       (N.FLAG_SYNC if sync else 0) | N.MODE_FLAGS[mode], None))
        return out

    def decode_sdf_jacobian(self, code, pts, mode: str = "tensor"):
        """(sdf (n,), d sdf/d[code,xyz] (n,67)) - loss_utils.py:82-103."""
        code = N.f32(code).reshape(64)
        pts = np.ascontiguousarray(N.f32(pts)[:, :3])
        n = pts.shape[0]
        sdf = np.empty(n, np.float32)
        jac = np.empty((n, N.JAC_STRIDE), np.float32)
        if n:
            N.check(self._L.qsp_decode_sdf_jacobian(self.handle, N.ptr(code), N.ptr(pts), n, N.ptr(sdf), N.ptr(jac),
                                                    N.FLAG_HOST_PTRS | N.MODE_FLAGS[mode], None))
        return sdf, jac[:, :67]

    def decode_sdf_jacobian_device(self, code, pts, sdf, jac, sync: bool = True, mode: str = "tensor"):
        N.check(self._L.qsp_decode_sdf_jacobian(self.handle, N.ptr(code), N.ptr(pts), pts.shape[0], N.ptr(sdf),
                                                N.ptr(jac), (N.FLAG_SYNC if sync else 0) | N.MODE_FLAGS[mode], None))

    def last_kernel_ms(self) -> float:
        ms = C.c_float()
        N.check(self._L.qsp_kernel_ms(self.handle, C.byref(ms)))
        return float(ms.value)


def as_device_decoder(decoder, device: Optional[int] = None) -> DeviceDecoder:
    if isinstance(decoder, DeviceDecoder):
        return decoder
    return DeviceDecoder(decoder, 0 if device is None else device)
