"""Per-iteration parity at the sizes that are benchmarked (BASELINE.json configs 1-3), product (tensor) mode, through the
batched C-ABI call: every object of one `kitti_batch32` step in ONE call (B = 32: the 9-chunk accumulation and the
single fwd+Jacobian launch of the batch path), `redwood_chair` (4 096 rays x 128 samples) and `kitti_single`
(2 200 rays x 50 samples; 64 chunks, side-stream surface rows).

The loop amplifies rounding differences, so parity is held PER ITERATION, from the oracle's own
trajectory: the fp32 oracle runs the 10 iterations; its state before iterations 0, 4 and 9 is handed to
the CUDA path for exactly one iteration and the two results are compared:
  * is_good identical;
  * in-ball sample count and the kept (ray, sample) list identical, in order - except entries whose deciding quantity
    (|x| - 1, |sdf| - th, de_do - 1e-2) lies within rounding of its threshold in the oracle's own numbers (SURVEY 8(d):
    such cases are counted and bounded, never silently accepted);
  * new pose <= 1e-3 for EVERY object-iteration (north_star); new code <= 1e-3, loss <= 1e-3, H <= 5e-4 and
    b <= 1e-3 (b relative to the magnitude of the sums it is the difference of; H and b without the gravity prior,
    whose 1 - cos(tilt) residual is compared on its own: it is a cancellation to the last fp32 digit multiplied by
    k4 = 1e7) for at least 90 % of them, with a hard bound of 2e-2 on every one; all relative to the largest entry
    (code: floor 0.05).
Why a quantile for the code: a ReLU pre-activation within rounding of zero flips a gradient row between ANY two fp32
implementations.  The reference's own fp32 and fp64 runs differ by 5e-4 .. 5e-3 on the code where that happens
(scripts/noise_floor.py, scripts/tmem_rz_emulation.py --full); the same spread is computed here for every case
(`ref_code`, fp32 oracle vs fp64 oracle) and written next to ours to gpurun_out/fullsize_parity*.jsonl.
"""
import json
import os

import numpy as np
import pytest
import torch

from qsp_slam_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHECK_ITERS = (0, 4, 9)


@pytest.fixture(scope="module")
def dev(decoder_state):
    from qsp_slam_b200.decoder import DeviceDecoder
    return DeviceDecoder(decoder_state, 0)


@pytest.fixture(scope="module")
def orc(decoder_state):
    from oracle import gn_oracle as G
    return G, G.FoldedDecoder(decoder_state)


def _rel(a, b, floor=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), floor))


def _borderline(ren, ray, sample, th):
    """Is (ray, sample) a case the oracle itself decides within rounding of a threshold?"""
    vi, vj = ren["valid_ray"].numpy(), ren["valid_sample"].numpy()
    k = np.nonzero((vi == ray) & (vj == sample))[0]
    if k.size == 0:
        return True                                        # not in the ball for the oracle: an in-ball borderline case
    s = float(ren["sdf"][k[0]])
    if abs(abs(s) - th) < 3e-5:
        return True
    bi, bj = ren["band_ray"].numpy(), ren["band_sample"].numpy()
    kb = np.nonzero((bi == ray) & (bj == sample))[0]
    if kb.size and abs(float(ren["de_do"][kb[0]]) - 1e-2) < 2e-4:
        return True
    return False


def _b_scale(G, prm, o):
    """Magnitude of the sums b is the (cancelling) difference of: k2/M |Js|^T |rs| + k1/P |Jr|^T |rr|."""
    st, rt = o["sdf"], o["render"]
    rs, _, _ = G.robust_res(st["res"], prm.b2)
    rr, _, _ = G.robust_res(rt["res"], prm.b1)
    Js = torch.cat([st["J_pose"], st["J_code"]], 1).abs()
    Jr = torch.cat([rt["J_pose"], rt["J_code"]], 1).abs()
    v = prm.k2 * (Js.t() @ rs.abs()) / Js.shape[0]
    if Jr.shape[0]:
        v = v + prm.k1 * (Jr.t() @ rr.abs()) / Jr.shape[0]
    return float(v.max())


# QSP_PARITY_MORE=1 (offline studies, several minutes of CPU oracle): more single-object cases of configs 1 and 3
_MORE = os.environ.get("QSP_PARITY_MORE", "0") != "0"


@pytest.mark.parametrize("workload,n_obj", [("kitti_batch32", 32), ("redwood_chair", 3 if _MORE else 1), ("kitti_single", 6 if _MORE else 1)])
def test_per_iteration_parity_at_benchmark_size(dev, decoder_state, orc, workload, n_obj):
    from qsp_slam_b200.reconstruct.optimizer import Optimizer
    from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
    from qsp_slam_b200 import _native as N
    G, fd = orc
    fd64 = G.FoldedDecoder(decoder_state, torch.float64)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cfg, objs = synth.make_workload(workload, decoder_state, objects=n_obj)
    prm = G.Params(cfg)
    th = prm.th
    # the oracle's own trajectories
    traces = []
    for o in objs:
        tr = []
        G.reconstruct_object(fd, prm, o["t_cam_obj"], o["pts"], o["rays"], o["depth"], trace=tr)
        traces.append(tr)
    cfg1 = json.loads(json.dumps(cfg))
    cfg1["optimizer"]["joint_optim"]["num_iterations"] = 1
    opt = Optimizer(dev, ForceKeyErrorDict(cfg1))
    stats, n_cases, n_thresh, n_samples = [], 0, 0, 0
    states = {}
    for e in CHECK_ITERS:
        idx = [i for i, tr in enumerate(traces) if len(tr) > e]            # objects the oracle carried this far
        assert len(idx) >= max(1, (3 * n_obj) // 4)
        dets, codes = [], []
        for i in idx:
            if e == 0:
                T, z = objs[i]["t_cam_obj"], np.zeros(64, np.float32)
            else:
                T = torch.inverse(traces[i][e - 1]["t_obj_cam"]).numpy()
                z = traces[i][e - 1]["code"].numpy()
            dets.append(dict(objs[i], t_cam_obj=np.ascontiguousarray(T, np.float32)))
            codes.append(np.ascontiguousarray(z, np.float32))
            t_oc = torch.inverse(torch.from_numpy(objs[i]["t_cam_obj"].copy())) if e == 0 else traces[i][e - 1]["t_obj_cam"]
            states[(i, e)] = (t_oc, torch.from_numpy(codes[-1].copy()), torch.from_numpy(np.ascontiguousarray(objs[i]["pts"])),
                              torch.from_numpy(np.ascontiguousarray(objs[i]["rays"])), torch.from_numpy(np.ascontiguousarray(objs[i]["depth"])))
        opt.set_debug(True)
        try:
            res = opt.reconstruct_objects(dets, codes)                     # ONE call for the whole batch
            its = [opt.get_iteration(0, b) for b in range(len(dets))]
            rots = []
            for b in range(len(dets)):                                     # [J_rot (7) | res_rot] of the recorded iteration
                rot = np.zeros(8, np.float32)
                N.check(dev._L.qsp_get_loss_rows(dev.handle, 0, b, 2, N.ptr(rot), 1))
                rots.append(rot)
        finally:
            opt.set_debug(False)
        for b, i in enumerate(idx):
            ref = traces[i][e]
            n_cases += 1
            assert res[b].is_good and not its[b]["failed"]
            ren = ref["render"]
            kr, ks = ren["keep_ray"].numpy(), ren["keep_sample"].numpy()
            n_samples += int(objs[i]["rays"].shape[0]) * prm.S
            d_ball = abs(its[b]["n_inball"] - ren["valid_ray"].shape[0])
            ours = set(zip(its[b]["keep_ray"].tolist(), its[b]["keep_sample"].tolist()))
            theirs = set(zip(kr.tolist(), ks.tolist()))
            diff = ours ^ theirs
            if not diff:
                assert np.array_equal(its[b]["keep_ray"], kr) and np.array_equal(its[b]["keep_sample"], ks)   # same ORDER
            # every differing entry must be a threshold case in the oracle's own numbers
            for (r_, s_) in diff:
                assert _borderline(ren, r_, s_, th), (workload, e, i, r_, s_)
            n_thresh += len(diff) + d_ball
            # the gravity prior (loss.py:155-178) is 1 - cos(tilt): near the upright pose the optimiser reaches, its
            # residual is a difference of two fp32 numbers that agree to 7 digits, and k4 = 1e7 multiplies it.  The
            # prior's own numbers are compared on their own, H and b without it.
            J_c, r_c = rots[b][:7].astype(np.float64), float(rots[b][7])
            J_o, r_o = ref["J_rot"].double().numpy(), float(ref["res_rot"])
            Hc = its[b]["H"].astype(np.float64)
            Hc[:7, :7] -= prm.k4 * np.outer(J_c, J_c)
            Ho = ref["H"].double().numpy().copy()
            Ho[:7, :7] -= prm.k4 * np.outer(J_o, J_o)
            bc = its[b]["b"].astype(np.float64)
            bc[:7] -= prm.k4 * J_c * r_c
            bo = ref["b"].double().numpy().copy()
            bo[:7] -= prm.k4 * J_o * r_o
            t_new = torch.inverse(ref["t_obj_cam"]).numpy()
            i64 = G.gn_iteration(fd64, prm, *(x.double() for x in states[(i, e)]))          # the oracle's own fp32 spread
            rec = dict(workload=workload, iteration=e, object=i, threshold_entries=len(diff) + d_ball,
                       H=_rel(Hc, Ho), b=float(np.abs(bc - bo).max() / _b_scale(G, prm, ref)),
                       rot_res=abs(r_c - r_o), rot_J=float(np.abs(J_c - J_o).max()),
                       dx=_rel(its[b]["dx"], ref["dx"].numpy()), pose=_rel(res[b].t_cam_obj, t_new),
                       code=_rel(res[b].code, ref["code"].numpy(), 0.05),
                       loss=abs(res[b].loss - float(ref["loss"])) / abs(float(ref["loss"])),
                       ref_code=_rel(ref["code"].numpy(), i64["code"].numpy(), 0.05) if i64["ok"] else None,
                       ref_pose=_rel(ref["t_obj_cam"].numpy(), i64["t_obj_cam"].numpy()) if i64["ok"] else None,
                       M=int(objs[i]["pts"].shape[0]), P=int(kr.shape[0]), Q=int(its[b]["n_inball"]))
            stats.append(rec)
    out = os.path.join(ROOT, "gpurun_out")
    q = lambda k, p: float(np.quantile([r[k] for r in stats if r[k] is not None], p))   # noqa: E731
    try:
        os.makedirs(out, exist_ok=True)
        keys = ("H", "b", "dx", "pose", "code", "loss", "rot_res", "rot_J", "ref_code", "ref_pose")
        summary = dict(workload=workload, cases=n_cases, threshold_entries=n_thresh, samples=n_samples,
                       max={k: q(k, 1.0) for k in keys}, p90={k: q(k, 0.9) for k in keys}, median={k: q(k, 0.5) for k in keys},
                       code_within_1e3=float(np.mean([r["code"] < 1e-3 for r in stats])))
        with open(os.path.join(out, "fullsize_parity.jsonl"), "a") as f:
            f.write(json.dumps(summary) + "\n")
        with open(os.path.join(out, "fullsize_parity_cases.jsonl"), "a") as f:
            for r in stats:
                f.write(json.dumps(r) + "\n")
    except OSError:
        pass
    # index sets: entries decided within rounding of a threshold are rare (and were each verified above)
    assert n_thresh <= 3 + 2e-5 * n_samples, (n_thresh, n_samples)
    for rec in stats:
        assert rec["pose"] < 1e-3, rec                      # north_star, every object-iteration
        assert rec["rot_res"] < 5e-7 and rec["rot_J"] < 1e-5, rec
        # hard bounds: a flipped ReLU row, a kept sample whose occupancy is within 1e-3 of 1 (de_do divides by 1 - o,
        # loss.py:122), or one threshold entry more or less among the P kept rows
        assert rec["code"] < 2e-2 and rec["H"] < 5e-2 and rec["b"] < 2e-2 and rec["loss"] < 1e-2, rec
    # north_star's 1e-3 on the code, and the H / b bounds of SURVEY 8(d), for at least 90 % of the object-iterations
    # (single objects: all but one of the three) - the reference's own fp32-vs-fp64 spread on the same cases
    # (ref_code) has the same tail
    need = len(stats) - max(1, len(stats) // 10)
    assert sum(r["code"] < 1e-3 for r in stats) >= need, sorted(r["code"] for r in stats)[-5:]
    assert sum(r["H"] < 5e-4 for r in stats) >= need, sorted(r["H"] for r in stats)[-5:]
    assert sum(r["b"] < 1e-3 for r in stats) >= need, sorted(r["b"] for r in stats)[-5:]
    assert sum(r["loss"] < 1e-3 for r in stats) >= need, sorted(r["loss"] for r in stats)[-5:]


def test_pose_only_at_benchmark_size(dev, decoder_state, orc):
    """Optimizer.estimate_pose_cam_obj (optimizer.py:47-93) at the size bench.py times it (`pose_only_objects_per_s`):
    the 32 objects of kitti_batch32, M = 2 000 surface points each, in ONE batched call, against the fp32 oracle run
    object by object - ONE iteration from the start pose (tight: the SDF term's 6 x 6 system is well conditioned), the
    config's five iterations, and eight (three iterations on the points kept by the inlier pruning, optimizer.py:80-82).  A start code of moderate norm
    keeps the surface off the zero-code sphere."""
    from qsp_slam_b200.reconstruct.optimizer import Optimizer
    from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict
    G, fd = orc
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    cfg, objs = synth.make_workload("kitti_batch32", decoder_state, objects=32)
    prm = G.Params(cfg)
    opt = Optimizer(dev, ForceKeyErrorDict(cfg))
    rng = np.random.default_rng(21)
    Ts, ss, codes = [], [], []
    for o in objs:
        T = o["t_cam_obj"].copy()
        s_ = np.cbrt(np.linalg.det(T[:3, :3].astype(np.float64)))
        T[:3, :3] /= s_
        Ts.append(np.ascontiguousarray(T, np.float32))
        ss.append(float(s_))
        codes.append((rng.standard_normal(64) * 0.03).astype(np.float32))
    pts = [o["pts"] for o in objs]
    worst = {}
    for iters, tol in ((1, 1e-4), (5, 1e-3), (8, 1e-3)):                      # 5 = the config's count; 8 runs past the pruning.  Observed: 1.1e-5, 3.1e-4, 2.9e-4
        opt.num_iterations_pose_only = iters
        prm.pose_only_iters = iters
        ours = opt.estimate_poses_cam_obj(Ts, ss, pts, codes)                 # one call for the batch
        errs = []
        for b in range(len(objs)):
            ref = G.estimate_pose_cam_obj(fd, prm, Ts[b], ss[b], pts[b], codes[b])
            errs.append(_rel(ours[b], ref))
        worst[iters] = max(errs)
        assert np.isfinite(ours).all() if isinstance(ours, np.ndarray) else all(np.isfinite(x).all() for x in ours)
        assert max(errs) < tol, (iters, max(errs))
        assert float(np.median(errs)) < tol / 5, (iters, float(np.median(errs)))
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "fullsize_parity.jsonl"), "a") as f:
        f.write(json.dumps({"workload": "pose_only_batch32", "max_rel_pose_error_by_iterations": worst}) + "\n")
