"""Drop-in for reconstruct/optimizer.py of the reference: same class names, constructor arguments,
method signatures, return object and failure convention; the work is done by the
sm_100a kernels behind include/qsp_recon.h.  There is no CPU path.

    Optimizer(decoder, configs).reconstruct_object(t_cam_obj, pts, rays, depth, code=None)
        -> ForceKeyErrorDict(t_cam_obj (4,4) f32 | None, code (64,) f32 | None, is_good, loss)
    Optimizer.estimate_pose_cam_obj(t_co_se3, scale, pts, code) -> (4,4)
    MeshExtractor(decoder, code_len, voxels_dim).extract_mesh_from_code(code) -> vertices, faces

Additions for throughput (not in the reference): ``reconstruct_objects`` (a batch of independent
detections in one launch sequence) and ``reconstruct_object_hypotheses`` (the flip_sample_num yaw
hypotheses of src/LocalMapping_util.cc:713-753 solved together, least loss among good ones wins).
"""
from __future__ import annotations

import ctypes as C
import math
import warnings
from typing import List, Optional, Sequence

import numpy as np

from .. import _native as N
from ..decoder import DeviceDecoder, as_device_decoder
from .utils import ForceKeyErrorDict, create_voxel_grid


def _cfg(obj, key):
    return obj[key] if isinstance(obj, dict) else getattr(obj, key)


class Optimizer(object):
    def __init__(self, decoder, configs, debug=False, device: Optional[int] = None):
        """reconstruct/optimizer.py:27-44.  ``decoder`` may be the deep_sdf ``Decoder`` module, its
        state dict, or a :class:`DeviceDecoder`."""
        self.decoder = as_device_decoder(decoder, device)
        optim_cfg = _cfg(configs, "optimizer")
        joint = _cfg(optim_cfg, "joint_optim")
        self.k1 = _cfg(joint, "k1")
        self.k2 = _cfg(joint, "k2")
        self.k3 = _cfg(joint, "k3")
        self.k4 = _cfg(joint, "k4")
        self.b1 = _cfg(joint, "b1")
        self.b2 = _cfg(joint, "b2")
        self.lr = _cfg(joint, "learning_rate")
        self.s_damp = _cfg(joint, "scale_damping")
        self.num_iterations_joint_optim = _cfg(joint, "num_iterations")
        self.code_len = _cfg(optim_cfg, "code_len")
        self.num_depth_samples = _cfg(optim_cfg, "num_depth_samples")
        self.cut_off = _cfg(optim_cfg, "cut_off_threshold")
        self.debug = debug
        if _cfg(configs, "data_type") == "KITTI":
            self.num_iterations_pose_only = _cfg(_cfg(optim_cfg, "pose_only_optim"), "num_iterations")
        if self.code_len != N.CODE_LEN:
            raise ValueError("code_len must be 64 for this decoder layout")

    # ------------------------------------------------------------------------------------------
    def _params(self) -> N.GnParams:
        return N.GnParams(self.k1, self.k2, self.k3, self.k4, self.b1, self.b2, self.lr, self.s_damp,
                          self.cut_off, int(self.num_iterations_joint_optim), int(self.num_depth_samples), 0)

    @staticmethod
    def _pack(arrs: Sequence[np.ndarray], width: int):
        """Concatenates per-object arrays into one C-ordered float32 array + (B+1) int32 offsets.  2-D inputs from
        pybind11/eigen.h are Fortran-ordered: they are copied column by column (contiguous reads), which
        is 3.5x faster than numpy's generic strided transpose copy (0.65 -> 0.18 ms for 32 x 2000 points)."""
        arrs = [np.asarray(a) for a in arrs]
        arrs = [a.reshape(-1, width) if width else a.reshape(-1) for a in arrs]
        off = np.zeros(len(arrs) + 1, np.int32)
        for i, a in enumerate(arrs):
            off[i + 1] = off[i] + a.shape[0]
        cat = np.empty((int(off[-1]), width) if width else (int(off[-1]),), np.float32)
        for i, a in enumerate(arrs):
            o = cat[off[i]:off[i + 1]]
            if not width or a.flags.c_contiguous:
                o[...] = a
            else:
                for c in range(width):
                    o[:, c] = a[:, c]
        return cat, off

    # ------------------------------------------------------------------------------------------
    def reconstruct_objects(self, detections: Sequence[dict], codes: Optional[Sequence] = None) -> List[ForceKeyErrorDict]:
        """Batch of independent detections; each is a mapping with ``t_cam_obj`` (4,4), ``pts`` (M,3),
        ``rays`` (N,3), ``depth`` (K,) (the fields of the reference's per-detection objects,
        reconstruct_frame.py:48-57).  One result per detection, in order."""
        if len(detections) == 0:
            return []
        rec = self.reconstruct_objects_records(detections, codes)
        return [self._result(rec[i]) for i in range(rec.shape[0])]

    def reconstruct_objects_records(self, detections: Sequence[dict], codes: Optional[Sequence] = None) -> np.ndarray:
        """Same solve, results as the raw (B,82) records (t_cam_obj 16 | code 64 | is_good | loss) that the
        multi-GPU all-gather moves (sharding.gather_records)."""
        B = len(detections)
        T = np.stack([N.f32(d["t_cam_obj"]).reshape(16) for d in detections])
        pts, pts_off = self._pack([d["pts"] for d in detections], 3)
        rays, rays_off = self._pack([d["rays"] for d in detections], 3)
        depth, depth_off = self._pack([d["depth"] for d in detections], 0)
        if depth.size == 0:
            depth = np.zeros(1, np.float32)
        z = None
        if codes is not None:
            z = np.stack([np.zeros(64, np.float32) if c is None else N.f32(c).reshape(-1)[:self.code_len]
                          for c in codes])
            z = np.ascontiguousarray(z, np.float32)
        rec = np.empty((B, N.RECORD_LEN), np.float32)
        prm = self._params()
        N.check(self.decoder._L.qsp_reconstruct_batch(self.decoder.handle, C.byref(prm), B, N.ptr(T), N.ptr(z),
                                              N.ptr(pts), N.i32p(pts_off), N.ptr(rays), N.i32p(rays_off),
                                              N.ptr(depth), N.i32p(depth_off), N.ptr(rec), N.FLAG_HOST_PTRS, None))
        self._check_call_stats()
        return rec

    def _check_call_stats(self):
        """Never raises (an exception would terminate the embedding SLAM process): a too-tight first-pass
        margin or exhausted candidate-row capacity is reported once per decoder as a warning; the affected objects
        of a capacity overflow already carry is_good = False."""
        try:
            st = self.decoder.call_stats()
        except Exception:
            return
        self.last_call_stats = st
        if st["n_margin_excursions"] > 0 and not getattr(self.decoder, "_warned_margin", False):
            self.decoder._warned_margin = True
            warnings.warn(f"qsp_slam_b200: the fp16 first pass was off by up to {st['max_first_pass_error']:.2e} on "
                          f"{st['n_margin_excursions']} re-evaluated samples (> 3/4 of the candidate margin): band samples "
                          "may be missed; call DeviceDecoder.calibrate_first_pass_margin() for this decoder", RuntimeWarning)
        if st["overflow"] and not getattr(self.decoder, "_warned_overflow", False):
            self.decoder._warned_overflow = True
            warnings.warn("qsp_slam_b200: candidate-row capacity exceeded; the objects that did not fit were returned "
                          "with is_good=False", RuntimeWarning)

    def pack(self, detections: Sequence[dict]):
        """Concatenated arrays + offsets of a detection list (the layout the C ABI takes)."""
        T = np.stack([N.f32(d["t_cam_obj"]).reshape(16) for d in detections])
        pts, pts_off = self._pack([d["pts"] for d in detections], 3)
        rays, rays_off = self._pack([d["rays"] for d in detections], 3)
        depth, depth_off = self._pack([d["depth"] for d in detections], 0)
        if depth.size == 0:
            depth = np.zeros(1, np.float32)
        return dict(T=T, pts=pts, pts_off=pts_off, rays=rays, rays_off=rays_off, depth=depth, depth_off=depth_off)

    def reconstruct_packed_device(self, packed: dict, records, codes=None, sync: bool = True, stream=None):
        """Same solve with inputs already resident in HBM: ``packed`` holds torch CUDA tensors for
        T/pts/rays/depth and host int32 numpy offsets; ``records`` (B,82) CUDA tensor is written."""
        B = packed["T"].shape[0]
        prm = self._params()
        N.check(self.decoder._L.qsp_reconstruct_batch(self.decoder.handle, C.byref(prm), B, N.ptr(packed["T"]), N.ptr(codes),
                                              N.ptr(packed["pts"]), N.i32p(packed["pts_off"]), N.ptr(packed["rays"]),
                                              N.i32p(packed["rays_off"]), N.ptr(packed["depth"]),
                                              N.i32p(packed["depth_off"]), N.ptr(records),
                                              N.FLAG_SYNC if sync else 0,
                                              None if stream is None else C.c_void_p(int(stream))))
        return records

    @staticmethod
    def _result(rec: np.ndarray) -> ForceKeyErrorDict:
        good = bool(rec[80] != 0)
        if not good:  # optimizer.py:162
            return ForceKeyErrorDict(t_cam_obj=None, code=None, is_good=False, loss=float(rec[81]))
        return ForceKeyErrorDict(t_cam_obj=rec[:16].reshape(4, 4).copy(), code=rec[16:80].copy(),
                                 is_good=True, loss=float(rec[81]))

    # KEY: the call of src/LocalMapping_util.cc:244-245 and :705-706
    def reconstruct_object(self, t_cam_obj, pts, rays, depth, code=None):
        """
        :param t_cam_obj: object pose, object-to-camera transformation (4, 4) Sim(3)
        :param pts: surface points, under camera coordinate (M, 3)
        :param rays: sampled ray directions (N, 3)
        :param depth: depth values (K,) only contain foreground pixels, K = M for KITTI
        :return: optimized object pose and shape, saved as a dict
        """
        det = dict(t_cam_obj=t_cam_obj, pts=pts, rays=rays, depth=depth)
        return self.reconstruct_objects([det], None if code is None else [code])[0]

    def reconstruct_object_hypotheses(self, t_cam_obj, pts, rays, depth, code=None, flip_sample_num: int = 4):
        """The orientation search of src/LocalMapping_util.cc:713-753 in one batch: hypothesis i rotates
        the object about its own y axis by i*2*pi/flip_sample_num; a later hypothesis replaces the
        current best if that one failed, or if it is good with a smaller loss.  Returns (best, all_results)."""
        T0 = N.f32(t_cam_obj).astype(np.float64)
        dets = []
        for i in range(flip_sample_num):
            a = i * 2.0 * math.pi / flip_sample_num
            Ry = np.array([[math.cos(a), 0, math.sin(a), 0], [0, 1, 0, 0], [-math.sin(a), 0, math.cos(a), 0],
                           [0, 0, 0, 1]])
            dets.append(dict(t_cam_obj=(T0 @ Ry).astype(np.float32), pts=pts, rays=rays, depth=depth))
        res = self.reconstruct_objects(dets, None if code is None else [code] * flip_sample_num)
        best = res[0]
        for r in res[1:]:  # the replacement rule of LocalMapping_util.cc:743-747, verbatim
            if (not best.is_good) or (best.loss > r.loss and r.is_good):
                best = r
        return best, res

    # the call of src/LocalMapping_util.cc:139-140
    def estimate_pose_cam_obj(self, t_co_se3, scale, pts, code):
        """
        :param t_co_se3: o2c transformation (4, 4) in SE(3)
        :param scale: object scale
        :param pts: surface points (M, 3)
        :param code: shape code
        :return: optimized o2c transformation
        """
        return self.estimate_poses_cam_obj([t_co_se3], [scale], [pts], [code])[0]

    def estimate_poses_cam_obj(self, t_co_se3s, scales, pts_list, codes):
        B = len(t_co_se3s)
        T = np.stack([N.f32(t).reshape(16) for t in t_co_se3s])
        sc = np.ascontiguousarray(np.asarray(scales, np.float32).reshape(B))
        z = np.ascontiguousarray(np.stack([N.f32(c).reshape(-1)[:64] for c in codes]), np.float32)
        pts, pts_off = self._pack([N.f32(p) for p in pts_list], 3)
        out = np.empty((B, 16), np.float32)
        N.check(self.decoder._L.qsp_estimate_pose_batch(self.decoder.handle, int(self.num_iterations_pose_only), B, N.ptr(T),
                                                N.ptr(sc), N.ptr(z), N.ptr(pts), N.i32p(pts_off), N.ptr(out),
                                                N.FLAG_HOST_PTRS, None))
        return [out[i].reshape(4, 4).copy() for i in range(B)]

    # ---- inspection for the parity tests --------------------------------------------------------
    def set_debug(self, on: bool):
        N.check(self.decoder._L.qsp_set_debug(self.decoder.handle, 1 if on else 0))

    def get_iteration(self, iteration: int, b: int = 0):
        H = np.empty((71, 71), np.float32)
        bv = np.empty(71, np.float32)
        dx = np.empty(71, np.float32)
        cnt = np.empty(6, np.int32)
        N.check(self.decoder._L.qsp_get_iteration(self.decoder.handle, iteration, b, N.ptr(H), N.ptr(bv), N.ptr(dx), N.ptr(cnt)))
        ray = np.empty(max(int(cnt[2]), 1), np.int32)
        smp = np.empty(max(int(cnt[2]), 1), np.int32)
        n = N.check(self.decoder._L.qsp_get_band_indices(self.decoder.handle, iteration, b, N.ptr(ray), N.ptr(smp), ray.size))
        n_cand = N.check(self.decoder._L.qsp_get_band_indices(self.decoder.handle, iteration, b, None, None, -1))
        return dict(H=H, b=bv, dx=dx, n_inball=int(cnt[0]), n_band=int(cnt[1]), n_keep=int(cnt[2]),
                    failed=bool(cnt[3]), keep_ray=ray[:n], keep_sample=smp[:n], n_cand=n_cand,
                    n_first_pass=int(cnt[4]))


class MeshExtractor(object):
    def __init__(self, decoder, code_len=64, voxels_dim=64, device: Optional[int] = None):
        """reconstruct/optimizer.py:284-290."""
        self.decoder = as_device_decoder(decoder, device)
        self.code_len = code_len
        self.voxels_dim = voxels_dim
        self.voxel_points = create_voxel_grid(vol_dim=self.voxels_dim)
        self._grid_on_device = False

    def _ensure_device_grid(self):
        """The grid points live on the device with the decoder handle (one handle may serve extractors of
        different voxels_dim: the last one set wins, so it is re-sent when another extractor replaced it)."""
        tag = (id(self), self.voxels_dim)
        if getattr(self.decoder, "_voxel_grid_tag", None) != tag:
            pts = np.ascontiguousarray(N.f32(self.voxel_points)[:, :3])
            N.check(self.decoder._L.qsp_set_voxel_grid(self.decoder.handle, N.ptr(pts), self.voxels_dim, N.FLAG_HOST_PTRS, None))
            self.decoder._voxel_grid_tag = tag

    def decode_grid(self, codes) -> np.ndarray:
        """SDF over the voxel grid for one code (64,) -> (D,D,D) or several (C,64) -> (C,D,D,D)
        (optimizer.py:296; the reference chunks by 64^3, loss_utils.py:51, the kernel does not need to)."""
        codes = N.f32(codes)
        single = codes.ndim == 1
        z = np.ascontiguousarray(codes.reshape(-1, codes.shape[-1])[:, :self.code_len])
        sdf = self.decoder.decode_sdf(z, self.voxel_points)
        D = self.voxels_dim
        sdf = sdf.reshape(-1, D, D, D)
        return sdf[0] if single else sdf

    def extract_mesh_from_code(self, code, exact_grid: bool = False, method: str = "cubes"):
        """optimizer.py:292-304: grid decode, then marching cubes at level 0 with spacing 2/(D-1) and
        origin (-1,-1,-1) (utils.py:120-141).  vertices (V,3) float32, faces (F,3) int32.  The vertices are those of
        the reference's Lewiner mesh (one per sign-changing grid edge); the triangulation comes from our own case
        table and is unpinned (scikit-image absent).  ``method="tets"``: marching tetrahedra.
        One native call (qsp_extract_mesh): the grid never leaves the device, and only the grid points whose value
        the isosurface depends on get the fp32-equivalent evaluation - the mesh is bit-identical to the one of the
        exact grid (``exact_grid=True``, the validation path)."""
        from ..mesh import extract_mesh
        self._ensure_device_grid()
        voxel_size = 2.0 / (self.voxels_dim - 1)
        code = N.f32(code).reshape(-1)[:self.code_len]
        vertices, faces = extract_mesh(self.decoder.handle, code, 0.0, voxel_size, (-1.0, -1.0, -1.0), exact_grid, method=method)
        return ForceKeyErrorDict(vertices=vertices, faces=faces)

    def extract_mesh_from_code_reference_grid(self, code, method: str = "cubes"):
        """The unfused form of the reference: ``decode_grid`` (fp32-equivalent everywhere, grid returned to the
        host) followed by the isosurface kernels on that host grid.  Kept for the tests."""
        from ..mesh import marching_cubes
        sdf = self.decode_grid(code)
        voxel_size = 2.0 / (self.voxels_dim - 1)
        vertices, faces = marching_cubes(sdf, 0.0, voxel_size, (-1.0, -1.0, -1.0), method=method)
        return ForceKeyErrorDict(vertices=vertices.astype("float32"), faces=faces.astype("int32"))

    def extract_meshes_from_codes(self, codes):
        """Batch form used by the offline driver (extract_map_objects.py:44-61): one native call per code, all on the
        device; nothing but the meshes comes back."""
        codes = np.asarray(codes, np.float32)
        return [self.extract_mesh_from_code(c) for c in codes.reshape(-1, codes.shape[-1])]
