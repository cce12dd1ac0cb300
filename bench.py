#!/usr/bin/env python
"""Benchmark of the object-reconstruction hot path (BASELINE.json metric: objects reconstructed/sec,
10 GN iterations; config: KITTI frame batch, 32 cars, 256 rays x 100 samples each, per GPU).

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA path through the C ABI)
  python bench.py --impl reference ...                     the CPU arm (oracle port of the reference)

One JSON line on rank 0.  A "step" is one pass of the hot path over one batch of synthetic objects
(32 per GPU: weak scaling, objects are independent; the only collective is the all-gather of the
82-float records).  See DESIGN.md "Measurement" for the meaning of every key.
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


def cpu_port_objects_per_s(state, cfg, objs, threads=None):
    """Times oracle.gn_oracle.reconstruct_object (the CPU port of the reference) on `objs`."""
    import torch
    from oracle import gn_oracle as G
    if threads:
        torch.set_num_threads(threads)
    fd = G.FoldedDecoder(state)
    prm = G.Params(cfg)
    t0 = time.perf_counter()
    for o in objs:
        G.reconstruct_object(fd, prm, o["t_cam_obj"], o["pts"], o["rays"], o["depth"])
    dt = time.perf_counter() - t0
    return len(objs) / dt, dt, torch.get_num_threads()


def run_reference(args, rank, real_stdout):
    """CPU arm: the oracle port with all host threads, each step a bounded sample of the workload."""
    if rank != 0:
        return
    from qsp_slam_b200 import synth
    import torch
    state = synth.make_decoder_state(0)
    n_sample = 8
    cfg, objs = synth.make_workload(WORKLOAD, state, objects=n_sample, pool_size=4096)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    for _ in range(min(args.warmup, 1)):
        cpu_port_objects_per_s(state, cfg, objs[:1])
    tot_t, tot_n = 0.0, 0
    for _ in range(args.steps):
        _, dt, th = cpu_port_objects_per_s(state, cfg, objs)
        tot_t += dt
        tot_n += len(objs)
    v = tot_n / tot_t
    sample = f"{n_sample} objects of {WORKLOAD} per step (of {OBJECTS_PER_GPU}), 10 GN iterations each"
    print(json.dumps({
        "impl": "reference", "metric": "objects reconstructed/sec (10 GN iters)", "value": v, "unit": "objects/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "objects_per_step": n_sample, "surface_pts": 2000, "rays": 256,
                   "samples": 100, "gn_iters": 10},
        "cpu_baseline": {"value": v, "unit": "objects/s", "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": v, "unit": "objects/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), file=real_stdout, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--objects", type=int, default=OBJECTS_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    state = synth.make_decoder_state(0)
    dev = DeviceDecoder(state, local_rank)
    decode = lambda c, p: dev.decode_sdf(c, p)  # noqa: E731  (input generation only)
    cfg, objs = synth.make_workload(WORKLOAD, state, objects=B, first_seed=1000 + 10000 * rank, decode=decode)
    opt = Optimizer(dev, ForceKeyErrorDict(cfg))
    packed = opt.pack(objs)
    h2d = sum(packed[k].nbytes for k in ("T", "pts", "rays", "depth", "pts_off", "rays_off", "depth_off"))
    d2h = B * 82 * 4
    dpk = {k: (torch.from_numpy(v).cuda() if k in ("T", "pts", "rays", "depth") else v) for k, v in packed.items()}
    rec = torch.empty((B, 82), dtype=torch.float32, device="cuda")
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def step():
        opt.reconstruct_packed_device(dpk, rec, sync=False)
        if world > 1:
            return sharding.gather_records(rec, [B] * world)
        return rec

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = _native.launch_count()
    evs = []
    for _ in range(args.steps):
        flush.zero_()                                  # L2 flush between timed steps (outside the events)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        step()
        e1.record()
        evs.append((e0, e1))
    barrier()
    launches = _native.launch_count() - l0
    ms = sum(a.elapsed_time(b) for a, b in evs)
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * B * args.steps / (ms_total / 1e3)
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
                              executed_tflops=n * ex * 1e6 / (msk * 1e-3) / 1e12)
        dom = max(kern, key=lambda k: kern[k]["ms"])
        traffic = None                                    # DRAM bytes per launch of the dominant kernel, from the committed ncu capture
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                traffic = json.load(f)[dom]["dram_bytes"]
        except Exception:
            pass
        roof = {"bound": "tensor", "kernel": dom, "achieved": kern[dom]["tflops"], "peak": pk["tflops"],
                "unit": "TFLOP/s", "frac": kern[dom]["tflops"] / pk["tflops"], "traffic": traffic,
                "peak_source": pk["source"] + " bf16 burst (kernel timed alone)",
                "algorithmic_flop_per_row": (FWD_MFLOP if dom == "decoder_fwd" else JAC_MFLOP) * 1e6,
                "executed": {"tflops": kern[dom]["executed_tflops"], "frac": kern[dom]["executed_tflops"] / pk["tflops"],
                             "note": "fp16 MMA flops issued by the kernel (fp32-equivalent arithmetic = 3 fp16 products)"},
                "kernels": kern}
        cpu = None
        if not args.no_cpu_baseline and world == 1:      # rank 0 at N=1 only (torchrun pins OMP_NUM_THREADS=1)
            ns = min(B, 32)
            v, dt, th = cpu_port_objects_per_s(state, cfg, objs[:ns])
            cpu = {"value": v, "unit": "objects/s", "cores": th, "kind": "port",
                   "sample": f"{ns} of the {B} objects of one step, 10 GN iterations each, oracle port, {dt:.1f} s"}
        out = {"metric": "objects reconstructed/sec (10 GN iters)", "value": value, "unit": "objects/s",
               "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
               "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": WORKLOAD, "objects_per_gpu": B, "surface_pts": 2000, "rays": 256,
                          "samples": 100, "gn_iters": 10, "decoder": "random-init calibrated DeepSDF 8x512, code 64",
                          "l2": "256 MB flush between timed steps", "parallelism": f"object-sharded x{world}",
                          "good_objects": good},
               "clocks": clocks, "gpu_launches": int(launches),
               "e2e": {"value": e2e, "unit": "objects/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
               "roofline": roof, "cpu_baseline": cpu,
               "decoder_evals_per_s": {k: v["rows"] / (v["ms"] * 1e-3) for k, v in kern.items()},
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
