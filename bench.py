#!/usr/bin/env python
"""Benchmark of the object-reconstruction hot path (BASELINE.json metric: objects reconstructed/sec,
10 GN iterations; config: KITTI frame batch, 32 cars, 256 rays x 100 samples each, per GPU).

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA path through the C ABI)
  python bench.py --impl reference ...                     the CPU arm: the UNMODIFIED reference (oracle/_ref,
                                                           byte-compiled from /root/reference by oracle/build_ref.py),
                                                           or the oracle port when that is absent

One JSON line on rank 0.  A "step" is one pass of the hot path over one batch of synthetic objects
(32 per GPU: weak scaling, objects are independent; the only collective is the all-gather of the
82-float records, issued on a side stream so that a rank never waits for a slower one inside a step).
See DESIGN.md "Measurement" for the meaning of every key.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "kitti_batch32"
OBJECTS_PER_GPU = 32
FWD_MFLOP = 3.67104      # 1 835 520 MAC per decoder evaluation
JAC_MFLOP = 7.34208      # forward + one reverse sweep
METRIC = "objects reconstructed/sec (10 GN iters)"
DTYPE = "fp16x2-split(3-product)/fp32-acc + fp16 first pass"
DECODER = "random-init calibrated DeepSDF 8x512, code 64"


def workload_config(world: int, B: int) -> dict:
    """The `config` object: the same for both arms."""
    return {"workload": WORKLOAD, "objects_per_gpu": B, "surface_pts": 2000, "rays": 256, "samples": 100,
            "gn_iters": 10, "decoder": DECODER, "l2": "256 MB flush between timed steps",
            "parallelism": f"object-sharded x{world}"}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(tflops=float(p["bf16_tflops"]), tflops_sustained=float(p["bf16_tflops_sustained"]),
                    hbm=float(p["hbm_gbs"]), source="measured")
    except Exception:
        return dict(tflops=1590.0, tflops_sustained=1400.0, hbm=6650.0, source="fallback")


class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "25", "-i", str(gpu)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        sm, mx, reasons = [], [], set()
        for line in open(self.f.name):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


# ------------------------------------------------------------------------------------------------------
# CPU arms (the only code of bench.py that touches oracle/)
# ------------------------------------------------------------------------------------------------------
class CpuArm:
    """The reference's CPU implementation of the path on the host cores: the UNMODIFIED reference when it can be
    imported (``oracle/ref_harness.py``: /root/reference here, oracle/_ref on the GPU box) - kind "reference" - else
    the oracle port - kind "port"."""

    def __init__(self, state, cfg, threads=None, force_port=False):
        import torch
        from oracle import gn_oracle as G
        from oracle import ref_harness as rh
        self.torch = torch
        self.G = G
        if threads:
            torch.set_num_threads(threads)
        self.threads = torch.get_num_threads()
        self.cfg = cfg
        self.kind = "port"
        self.fd = G.FoldedDecoder(state)
        self.prm = G.Params(cfg)
        self.rh = rh
        if rh.available() and not force_port:
            try:
                ref = rh.load()
                self.opt = ref.optimizer.Optimizer(rh.build_decoder(state), rh.make_configs(cfg))
                self.kind = "reference"
            except Exception as e:  # a broken oracle/_ref must not take the bench line down
                print(f"bench: reference import failed ({e!r}); timing the oracle port", file=sys.stderr)
            finally:
                rh.restore_cuda()       # the harness patches Tensor.cuda process-wide: only while the reference runs

    def run(self, objs, kind=None):
        """-> (results [dict(is_good, t_cam_obj, code, loss)], seconds)"""
        kind = kind or self.kind
        out = []
        t0 = time.perf_counter()
        if kind == "reference":
            with self.rh.cpu_only():    # .cuda() -> identity only inside this block (the GPU arm shares the process)
                for o in objs:
                    r = self.opt.reconstruct_object(o["t_cam_obj"].copy(), o["pts"], o["rays"], o["depth"])
                    out.append(dict(is_good=bool(r.is_good), t_cam_obj=r.t_cam_obj, code=r.code, loss=float(r.loss)))
        else:
            for o in objs:
                out.append(self.G.reconstruct_object(self.fd, self.prm, o["t_cam_obj"], o["pts"], o["rays"], o["depth"]))
        return out, time.perf_counter() - t0

    def first_iteration(self, o):
        """index sets of iteration 0 from the oracle port (the reference does not expose them)"""
        t = self.torch
        return self.G.gn_iteration(self.fd, self.prm, t.inverse(t.from_numpy(o["t_cam_obj"].copy())), t.zeros(64),
                                   t.from_numpy(np.ascontiguousarray(o["pts"])), t.from_numpy(np.ascontiguousarray(o["rays"])),
                                   t.from_numpy(np.ascontiguousarray(o["depth"])))


def run_reference(args, rank, real_stdout):
    """CPU arm: same workload and config as ours; every step is a bounded sample of the step's 32 objects (a window
    that rotates through all of them), sized from a warm-up rate so that the whole run ends within a few minutes."""
    if rank != 0:
        return
    from qsp_slam_b200 import synth
    state = synth.make_decoder_state(0)
    B = args.objects
    cfg, objs = synth.make_workload(WORKLOAD, state, objects=B, pool_size=4096)
    cores = os.cpu_count() or 1
    arm = CpuArm(state, cfg, threads=cores)
    _, dt = arm.run(objs[:1])
    _, dt = arm.run(objs[1:3])                                   # warm-up and rate estimate
    rate = 2.0 / dt
    budget_s = 200.0
    n_step = int(max(4, min(B, budget_s * rate / max(args.steps, 1))))
    tot_t, tot_n, pos = 0.0, 0, 0
    for _ in range(args.steps):
        sel = [objs[(pos + i) % B] for i in range(n_step)]
        pos = (pos + n_step) % B
        _, dt = arm.run(sel)
        tot_t += dt
        tot_n += n_step
    v = tot_n / tot_t
    sample = (f"{n_step} of the {B} objects of one {WORKLOAD} step per bench step (rotating window over all {B}), "
              f"10 GN iterations each, {'unmodified reference (oracle/_ref)' if arm.kind == 'reference' else 'oracle port'}, "
              f"{tot_t:.1f} s in all")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "objects/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(args.steps, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus, B),
        "cpu_baseline": {"value": v, "unit": "objects/s", "cores": arm.threads, "kind": arm.kind, "sample": sample},
        "e2e": {"value": v, "unit": "objects/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), file=real_stdout, flush=True)


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
def other_configs(torch, dev, state, decode, pk):
    """The other BASELINE.json configs, each timed on its own outside the headline's timed region (median of 5,
    host buffers in, host results out unless stated).  Parity for these sizes: tests/test_gpu_fullsize_parity.py."""
    from qsp_slam_b200 import synth
    from qsp_slam_b200.reconstruct.optimizer import Optimizer, MeshExtractor
    from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict, create_voxel_grid

    def timed(fn, reps=5, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        return float(np.median(ts))

    out = {}
    for name in ("kitti_single", "redwood_chair"):
        cfg, objs = synth.make_workload(name, state, decode=decode)
        opt = Optimizer(dev, ForceKeyErrorDict(cfg))
        o = objs[0]
        out[name + "_ms"] = 1e3 * timed(lambda: opt.reconstruct_object(o["t_cam_obj"], o["pts"], o["rays"], o["depth"]))
        if name == "redwood_chair":
            out["hypotheses4_ms"] = 1e3 * timed(lambda: opt.reconstruct_object_hypotheses(
                o["t_cam_obj"], o["pts"], o["rays"], o["depth"], flip_sample_num=4))
    cfg, objs = synth.make_workload(WORKLOAD, state, decode=decode)
    opt = Optimizer(dev, ForceKeyErrorDict(cfg))
    Ts, ss = [], []
    for o in objs:
        T = o["t_cam_obj"].copy()
        s = np.cbrt(np.linalg.det(T[:3, :3].astype(np.float64)))
        T[:3, :3] /= s
        Ts.append(T)
        ss.append(float(s))
    z = [np.zeros(64, np.float32)] * len(objs)
    tp = timed(lambda: opt.estimate_poses_cam_obj(Ts, ss, [o["pts"] for o in objs], z))
    out["pose_only_objects_per_s"] = len(objs) / tp
    # config 4: 64 codes x 128^3 grid points, fp32-equivalent decode, device resident (kernel time, CUDA events)
    grid = torch.from_numpy(create_voxel_grid(128)).cuda()
    codes = (torch.randn(64, 64, generator=torch.Generator().manual_seed(0)) * 0.1).cuda()
    res = torch.empty((64, grid.shape[0]), device="cuda")
    for _ in range(2):
        dev.decode_sdf_device(codes, grid, res, mode="tensor")
    s_grid = dev.last_kernel_ms() / 1e3
    n = 64 * grid.shape[0]
    tf = n * FWD_MFLOP * 1e6 / s_grid / 1e12
    out["mesh_grid_128x64"] = {"s": s_grid, "tflops_algorithmic": tf, "frac": tf / pk["tflops_sustained"],
                               "peak": "measured bf16 sustained" if pk["source"] == "measured" else "fallback sustained",
                               "evals_per_s": n / s_grid}
    del res
    mx = MeshExtractor(dev, 64, 128)
    cz = codes.cpu().numpy()
    mx.extract_mesh_from_code(cz[0])
    out["mesh_extract_128_ms"] = 1e3 * timed(lambda: [mx.extract_mesh_from_code(c) for c in cz[:8]], reps=3, warm=1) / 8
    return out


def parity_block(opt, objs, cpu_arm, cpu_results, gpu_records):
    """Parity evidence on the benchmarked objects themselves (N = 1): is_good of the full loop against the CPU arm's
    results, and the index sets of iteration 0 (in-ball count, kept (ray, sample) list in order) against the oracle."""
    B = len(objs)
    opt.set_debug(True)
    try:
        saved = opt.num_iterations_joint_optim
        opt.num_iterations_joint_optim = 1
        opt.reconstruct_objects(objs)
        its = [opt.get_iteration(0, b) for b in range(B)]
    finally:
        opt.num_iterations_joint_optim = saved
        opt.set_debug(False)
    n_inball = n_keep = 0
    for o, it in zip(objs, its):
        ref = cpu_arm.first_iteration(o)
        if not ref["ok"]:
            continue
        ren = ref["render"]
        n_inball += int(it["n_inball"] == ren["valid_ray"].shape[0])
        n_keep += int(np.array_equal(it["keep_ray"], ren["keep_ray"].numpy()) and
                      np.array_equal(it["keep_sample"], ren["keep_sample"].numpy()))
    out = {"objects": B, "iter0_inball_count_equal": n_inball, "iter0_kept_ray_sample_lists_equal": n_keep,
           "iter0_checked_against": "oracle port (fp32, same inputs)"}
    if cpu_results is not None:
        m = len(cpu_results)
        good_gpu = [bool(gpu_records[i, 80] != 0) for i in range(m)]
        out["is_good_equal"] = int(sum(int(g == bool(r["is_good"])) for g, r in zip(good_gpu, cpu_results)))
        out["is_good_compared"] = m
        dp, dc = [], []
        for i, r in enumerate(cpu_results):
            if r["is_good"] and good_gpu[i]:
                t = np.asarray(r["t_cam_obj"], np.float64).reshape(16)
                dp.append(float(np.abs(gpu_records[i, :16] - t).max() / np.abs(t).max()))
                c = np.asarray(r["code"], np.float64)
                dc.append(float(np.abs(gpu_records[i, 16:80] - c).max() / max(np.abs(c).max(), 0.05)))
        if dp:
            out["final_pose_rel_diff_median"] = float(np.median(dp))
            out["final_code_rel_diff_median"] = float(np.median(dc))
            out["final_note"] = "10 chained iterations: the loop amplifies rounding differences (DESIGN.md section 4); " \
                                "the per-iteration bounds are in tests/test_gpu_fullsize_parity.py"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--objects", type=int, default=OBJECTS_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the other BASELINE configs (quick runs)")
    ap.add_argument("--map-objects", type=int, default=1024, help="N > 1: objects of the strong-scaling map batch (0 = skip)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    real_stdout = _json_only_stdout()
    if args.impl == "reference":
        run_reference(args, rank, real_stdout)
        return

    import torch
    import torch.distributed as dist
    from qsp_slam_b200 import synth, _native, sharding
    from qsp_slam_b200.decoder import DeviceDecoder
    from qsp_slam_b200.reconstruct.optimizer import Optimizer
    from qsp_slam_b200.reconstruct.utils import ForceKeyErrorDict

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    B = args.objects
    W = max(args.warmup, 3)
    state = synth.make_decoder_state(0)
    dev = DeviceDecoder(state, local_rank)
    decode = lambda c, p: dev.decode_sdf(c, p)  # noqa: E731  (input generation only)
    # the job's objects: world * B distinct synthetic cars, dealt to the ranks by a host-side cost proxy so that the
    # per-rank work is balanced; every rank generates the same list (seeded) and takes its share
    cfg, all_objs = synth.make_workload(WORKLOAD, state, objects=world * B, first_seed=1000, decode=decode)
    S = cfg["optimizer"]["num_depth_samples"]
    proxy = [sharding.inball_cost(o, S) for o in all_objs]
    shares = sharding.refine_equal_counts(proxy, sharding.balanced_equal_counts(proxy, world))
    objs = [all_objs[i] for i in shares[rank]]
    opt = Optimizer(dev, ForceKeyErrorDict(cfg))
    balance = "equal counts by a host proxy (in-ball samples at the initial pose), refined by swaps"
    packed = opt.pack(objs)
    h2d = sum(packed[k].nbytes for k in ("T", "pts", "rays", "depth", "pts_off", "rays_off", "depth_off"))
    d2h = B * 82 * 4
    dpk = {k: (torch.from_numpy(v).cuda() if k in ("T", "pts", "rays", "depth") else v) for k, v in packed.items()}
    recs = [torch.empty((B, 82), dtype=torch.float32, device="cuda") for _ in range(2)]   # double buffered
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2
    gatherer = sharding.RecordGatherer([B] * world, "cuda") if world > 1 else None
    side = torch.cuda.Stream() if world > 1 else None
    main_stream = torch.cuda.current_stream()

    def run_steps(n):
        """n steps back to back.  Step k: L2 flush, the GN solve of the rank's batch (async), then - on a side stream,
        behind an event - the all-gather of its records: the next step's kernels never wait for a slower rank.
        Returns per-step (compute start, compute end, gather end) events."""
        evs = []
        for k in range(n):
            rec = recs[k & 1]
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            opt.reconstruct_packed_device(dpk, rec, sync=False)
            e1.record()
            e2 = None
            if world > 1:
                side.wait_event(e1)
                with torch.cuda.stream(side):
                    gatherer.gather(rec)
                    e2 = torch.cuda.Event(enable_timing=True)
                    e2.record()
                if k >= 1:
                    main_stream.wait_event(evs[k - 1][2])      # rec[k & 1] is reused two steps later
            evs.append((e0, e1, e2))
        if world > 1:
            main_stream.wait_stream(side)
        return evs

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    run_steps(W)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = _native.launch_count()
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin.record()
    evs = run_steps(args.steps)
    t_end.record()
    barrier()
    launches = _native.launch_count() - l0
    ms = t_begin.elapsed_time(t_end)                   # the whole timed region, flushes and gathers included
    compute_ms = sum(a.elapsed_time(b) for a, b, _ in evs)
    gather_ms = sum(b.elapsed_time(c) for _, b, c in evs) if world > 1 else 0.0
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms, compute_ms, gather_ms], dtype=torch.float64, device="cuda")
    tmax, tmin = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    ms_total = float(tmax[0].item())
    value = world * B * args.steps / (ms_total / 1e3)
    rec = recs[(args.steps - 1) & 1]
    good = int((rec[:, 80] != 0).sum().item())

    # end to end through the public API with host buffers (h2d + d2h inside the timed region)
    opt.reconstruct_objects(objs)
    barrier()
    t0 = time.perf_counter()
    n_e2e = max(2, min(args.steps, 5))
    for _ in range(n_e2e):
        res = opt.reconstruct_objects(objs)
    torch.cuda.synchronize()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e = world * B * n_e2e / float(te.item())

    # per rank: rounds of the dominant kernel summed over the iterations of one step (k_decoder_p3 walks
    # ceil(tiles / SMs) rounds of equal length: a share with slightly more candidate rows crosses into another round)
    opt.set_debug(True)
    opt.reconstruct_objects(objs)
    n_it = cfg["optimizer"]["joint_optim"]["num_iterations"]
    m_rows = sum(o["pts"].shape[0] for o in objs)
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    jac_rows = [m_rows + sum(opt.get_iteration(i, b)["n_cand"] for b in range(B)) for i in range(n_it)]
    opt.set_debug(False)
    jac_rounds = sum(-(-(-(-r // 64)) // sms) for r in jac_rows)
    tr = torch.tensor([float(jac_rounds), float(sum(jac_rows))], dtype=torch.float64, device="cuda")
    trmax, trmin = tr.clone(), tr.clone()
    if world > 1:
        dist.all_reduce(trmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(trmin, op=dist.ReduceOp.MIN)

    # BASELINE config 5 next to the weak-scaling headline: a fixed map of `map_objects` objects, STRONG scaling
    map_scale = None
    if world > 1 and args.map_objects > 0:
        n_map = args.map_objects
        distinct = min(64, world * B)
        base = all_objs[:distinct]
        mobjs = [base[i % distinct] for i in range(n_map)]
        lo, hi = sharding.partition(n_map, world)[rank]
        mine = mobjs[lo:hi]
        counts = [h - l for l, h in sharding.partition(n_map, world)]
        g2 = sharding.RecordGatherer(counts, "cuda")

        def run_map():
            parts = [opt.reconstruct_objects_records(mine[b0:b0 + 128]) for b0 in range(0, len(mine), 128)]
            loc = torch.from_numpy(np.concatenate(parts, 0)).cuda()
            g2.gather(loc)
            return g2.ordered()

        run_map()
        barrier()
        m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        m0.record()
        allrec = run_map()
        m1.record()
        torch.cuda.synchronize()
        mt = torch.tensor([m0.elapsed_time(m1)], dtype=torch.float64, device="cuda")
        dist.all_reduce(mt, op=dist.ReduceOp.MAX)
        map_scale = {"workload": "map1024", "objects": n_map, "scaling": "strong", "seconds": float(mt.item()) / 1e3,
                     "objects_per_s": n_map / (float(mt.item()) / 1e3), "good_objects": int((allrec[:, 80] != 0).sum().item()),
                     "api": "Optimizer.reconstruct_objects (host buffers, sub-batches of 128) + all_gather of the records"}

    out = None
    if rank == 0:
        pk = peaks()
        # dominant kernels timed alone on the row counts of a real step (CUDA events on the launch stream)
        opt.set_debug(True)
        opt.reconstruct_objects(objs)
        its = [opt.get_iteration(5, b) for b in range(B)]
        opt.set_debug(False)
        q_rows = sum(i["n_first_pass"] for i in its)   # ray samples the first pass decodes (of n_inball in the ball)
        j_rows = sum(o["pts"].shape[0] for o in objs) + sum(i["n_cand"] for i in its)
        g = torch.Generator(device="cuda").manual_seed(1)

        def ball(n):
            x = torch.randn((n, 3), generator=g, device="cuda")
            return (x / x.norm(dim=1, keepdim=True) * torch.rand((n, 1), generator=g, device="cuda") ** (1 / 3)).contiguous()

        zc = torch.zeros(64, device="cuda")
        kern = {}
        for name, n, mf in (("decoder_fwd", q_rows, FWD_MFLOP), ("decoder_fwd_jac", j_rows, JAC_MFLOP)):
            # decoder_fwd: single-pass fp16 tcgen05 kernel (1 MMA pass per layer);
            # decoder_fwd_jac: fp16-pair tcgen05 kernel (3 MMA products per algorithmic layer pass)
            x = ball(n)
            so = torch.empty(n, device="cuda")
            jo = torch.empty((n, 68), device="cuda")
            tms = []
            for i in range(6):
                flush.zero_()
                if name == "decoder_fwd":
                    dev.decode_sdf_device(zc, x, so, mode="fast")       # first pass over the ray samples
                else:
                    dev.decode_sdf_jacobian_device(zc, x, so, jo)       # fp16-pair kernel, fwd + reverse sweep
                if i >= 3:
                    tms.append(dev.last_kernel_ms())
            msk = float(np.mean(tms))
            # MMA flops the kernel actually issues per row: zero-padded layer 0 (K = 128), and for the fp16-pair
            # kernel 3 partial products per contraction (k_decoder_p3) plus the N = 256 last reverse step
            ex = 3.801088 if name == "decoder_fwd" else 3 * (3.801088 + 3.670016 + 0.262144)
            kern[name] = dict(rows=n, ms=msk, tflops=n * mf * 1e6 / (msk * 1e-3) / 1e12,
                              executed_tflops=n * ex * 1e6 / (msk * 1e-3) / 1e12, evals_per_s=n / (msk * 1e-3))
        dom = max(kern, key=lambda k: kern[k]["ms"])
        traffic = None                                    # DRAM bytes per launch of the dominant kernel, from the committed ncu capture
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                traffic = json.load(f)[dom]["dram_bytes"]
        except Exception:
            pass
        mf_dom = FWD_MFLOP if dom == "decoder_fwd" else JAC_MFLOP
        roof_evals = pk["tflops_sustained"] * 1e12 / (mf_dom * 1e6)       # 190 M (fwd+Jacobian)
        roof = {"bound": "tensor", "kernel": dom, "achieved": kern[dom]["tflops"], "peak": pk["tflops"],
                "unit": "TFLOP/s", "frac": kern[dom]["tflops"] / pk["tflops"], "traffic": traffic,
                "peak_source": pk["source"] + " bf16 burst (kernel timed alone)",
                "algorithmic_flop_per_row": mf_dom * 1e6,
                "evals_per_s": kern[dom]["evals_per_s"],
                "evals_roofline": {"per_s": roof_evals, "frac": kern[dom]["evals_per_s"] / roof_evals,
                                   "target_60pct_per_s": 0.6 * roof_evals,
                                   "peak": pk["source"] + " bf16 sustained (BASELINE.md section 2)"},
                "executed": {"tflops": kern[dom]["executed_tflops"], "frac": kern[dom]["executed_tflops"] / pk["tflops"],
                             "note": "fp16 MMA flops issued by the kernel (fp32-equivalent arithmetic = 3 fp16 products)"},
                "kernels": kern}
        cpu = parity = None
        if not args.no_cpu_baseline and world == 1:      # rank 0 at N=1 only (torchrun pins OMP_NUM_THREADS=1)
            arm = CpuArm(state, cfg)
            # bounded samples (~10-30 s of CPU work in all): the unmodified reference on the first 8 objects of the
            # step (it is ~4x slower than the port: autograd, weight norm re-evaluated per call), the port on all
            # of them - the port's results also feed the parity block
            ns = min(B, 8) if arm.kind == "reference" else min(B, 32)
            cres, dt = arm.run(objs[:ns])
            cpu = {"value": ns / dt, "unit": "objects/s", "cores": arm.threads, "kind": arm.kind,
                   "sample": f"{ns} of the {B} objects of one step, 10 GN iterations each, "
                             f"{'unmodified reference (oracle/_ref)' if arm.kind == 'reference' else 'oracle port'}, {dt:.1f} s"}
            if arm.kind == "reference":
                pres, dtp = arm.run(objs[:min(B, 32)], kind="port")
                cpu["port_value"] = min(B, 32) / dtp
                cpu["reference_vs_port_is_good_equal"] = int(sum(int(a["is_good"] == b["is_good"]) for a, b in zip(cres, pres)))
                cres = pres
            parity = parity_block(opt, objs, arm, cres, rec.cpu().numpy())
        configs = None
        if not args.no_configs and world == 1:
            configs = other_configs(torch, dev, state, decode, pk)
        cfgd = workload_config(world, B)                 # identical to the reference arm's `config` (same_config)
        out = {"metric": METRIC, "value": value, "unit": "objects/s",
               "n_gpus": world, "steps": args.steps, "warmup": W,
               "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
               "config": cfgd, "good_objects": good, "shares": balance,
               "clocks": clocks, "gpu_launches": int(launches),
               "e2e": {"value": e2e, "unit": "objects/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
               "roofline": roof, "cpu_baseline": cpu, "parity": parity, "configs": configs,
               "ranks": {"timed_region_ms": {"max": float(tmax[0]), "min": float(tmin[0])},
                         "compute_ms_sum": {"max": float(tmax[1]), "min": float(tmin[1])},
                         "gather_tail_ms_sum": {"max": float(tmax[2]), "min": float(tmin[2])},
                         "fwd_jac_rounds_per_step": {"max": int(trmax[0]), "min": int(trmin[0])},
                         "fwd_jac_rows_per_step": {"max": int(trmax[1]), "min": int(trmin[1])},
                         "note": "per rank: whole timed region; sum over steps of the GN solve; sum over steps of "
                                 "(gather end - solve end) on the side stream; rounds (ceil(tiles / SMs), summed over the "
                                 "iterations of one step) and rows of the fwd+Jacobian launch"},
               "map_scale": map_scale,
               "decoder_evals_per_s": {k: v["evals_per_s"] for k, v in kern.items()},
               "rows_per_iteration": {"in_ball_samples": sum(i["n_inball"] for i in its), "first_pass": q_rows,
                                      "fwd_jac": j_rows, "kept_band": sum(i["n_keep"] for i in its)}}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), file=real_stdout, flush=True)


def _json_only_stdout():
    """Everything libraries write to file descriptor 1 (NCCL prints its version banner there) goes to stderr; the
    returned file is the real stdout, used for the one JSON line."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    main()
