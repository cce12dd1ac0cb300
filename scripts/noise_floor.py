"""TEST INFRASTRUCTURE (build container only: imports the unmodified reference through oracle/ref_harness.py).

What one Gauss-Newton iteration (reconstruct/optimizer.py:141-263) can be held to: the spread of the reference's OWN
arithmetic.  For every fixture object of tests/golden (and, with --full, one object of each BASELINE size) one
iteration is run from (t_cam_obj, code0) by
  ref1   the unmodified reference, 1 CPU thread            (the configuration of record)
  refN   the unmodified reference, all CPU threads
  o32    oracle/gn_oracle.py in fp32
  o64    oracle/gn_oracle.py in fp64                       (taken as the truth)
  emu3   the oracle with the decoder's contractions carried as three fp16-pair products
         (x_hi w_hi + x_lo w_hi + x_hi w_lo, exact accumulation): the arithmetic of k_decoder_p3 without the
         tensor pipe's fp32 accumulation error - what the operand split alone costs
and the new pose / new code / b / dx are compared, relative to the largest entry (code: floor 0.05, as the tests do).
Writes a markdown table to stdout; DESIGN.md section 4 quotes it.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from qsp_slam_b200 import synth
from oracle import gn_oracle as G
from oracle import ref_harness as rh


def rel(a, b, floor=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), floor))


def hi(t):
    return t.half().double()


def lo(t):
    return (t - hi(t)).half().double()


class Emu3(G.FoldedDecoder):
    """fp64 container whose matmuls are evaluated as three fp16-pair partial products of the fp32 operands."""
    pass


def emu_mm(a, w):
    a = a.float().double()
    w = w.float().double()
    ah, al, wh, wl = hi(a * 16.0), lo(a * 16.0), hi(w * 16.0), lo(w * 16.0)
    return (ah @ wh.t() + al @ wh.t() + ah @ wl.t()) / 256.0


def emu_forward_jacobian(dec, code, x):
    n = x.shape[0]
    inp = torch.cat([code.double().expand(n, -1), x.double()], 1)
    a = inp
    masks = []
    for l in range(8):
        if l == 4:
            a = torch.cat([a, inp], 1)
        a = (emu_mm(a, dec.W[l]) + dec.b[l]).float().double()
        masks.append(a > 0)
        a = torch.relu(a)
    s = torch.tanh((a @ dec.W[8][0] + dec.b[8][0]).float().double())
    g = ((1 - s * s)[:, None] * dec.W[8]) * masks[7]
    skip = None
    for l in range(7, 0, -1):
        g = emu_mm(g * 256.0, dec.W[l].t()) / 256.0
        if l == 4:
            skip = g[:, -67:]
            g = g[:, :-67]
        g = (g * masks[l - 1]).float().double()
    g = emu_mm(g * 256.0, dec.W[0].t()) / 256.0 + skip
    return s, g.float().double()


def one_iteration_oracle(fd, prm, o, code0, emu=False):
    dt = fd.dtype
    t_oc = torch.inverse(torch.from_numpy(o["t_cam_obj"].copy()).to(dt))
    args = (fd, prm, t_oc, torch.from_numpy(code0.copy()).to(dt), torch.from_numpy(np.ascontiguousarray(o["pts"])).to(dt),
            torch.from_numpy(np.ascontiguousarray(o["rays"])).to(dt), torch.from_numpy(np.ascontiguousarray(o["depth"])).to(dt))
    if not emu:
        return G.gn_iteration(*args)
    keep = G.decoder_forward_jacobian
    G.decoder_forward_jacobian = emu_forward_jacobian
    try:
        return G.gn_iteration(*args)
    finally:
        G.decoder_forward_jacobian = keep


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true", help="also one object of each BASELINE size (slow)")
    args = ap.parse_args()
    ref = rh.load()
    state = synth.make_decoder_state(0)
    dec = rh.build_decoder(state)
    golden = dict(np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "reference_outputs.npz")))
    cases = []
    for name, style in (("kitti0", "kitti"), ("kitti1", "kitti"), ("redwood0", "redwood"), ("redwood1", "redwood")):
        p = name + "_"
        o = {k: golden[p + k] for k in ("t_cam_obj", "pts", "rays", "depth")}
        cases.append((name, style, 50, o, golden[p + "code0"]))
    if args.full:
        for wl, style in (("kitti_single", "kitti"), ("kitti_batch32", "kitti"), ("redwood_chair", "redwood")):
            cfg, objs = synth.make_workload(wl, state, objects=1, pool_size=16384)
            code0 = (np.random.default_rng(100).standard_normal(64) * 0.03).astype(np.float32)
            cases.append((wl, style, cfg["optimizer"]["num_depth_samples"], objs[0], code0))
    nthreads = os.cpu_count() or 1
    fd32, fd64 = G.FoldedDecoder(state), G.FoldedDecoder(state, torch.float64)
    print("| object | quantity | ref1 vs o64 | refN vs ref1 | o32 vs o64 | o32 vs ref1 | emu3 vs o64 |")
    print("|---|---|---|---|---|---|---|")
    for name, style, S, o, code0 in cases:
        cfg = synth.optimizer_config(style, num_depth_samples=S, num_iterations=1)
        prm = G.Params(cfg)
        outs = {}
        for tag, nt in (("ref1", 1), ("refN", nthreads)):
            torch.set_num_threads(nt)
            opt = ref.optimizer.Optimizer(dec, rh.make_configs(cfg))
            r = opt.reconstruct_object(o["t_cam_obj"].copy(), o["pts"], o["rays"], o["depth"], code0.copy())
            outs[tag] = dict(t=np.linalg.inv(r.t_cam_obj.astype(np.float64)), code=r.code)
        torch.set_num_threads(1)
        i32 = one_iteration_oracle(fd32, prm, o, code0)
        torch.set_num_threads(nthreads)
        i64 = one_iteration_oracle(fd64, prm, o, code0)
        ie = one_iteration_oracle(fd64, prm, o, code0, emu=True)
        for tag, it in (("o32", i32), ("o64", i64), ("emu3", ie)):
            outs[tag] = dict(t=it["t_obj_cam"].double().numpy(), code=it["code"].double().numpy(), b=it["b"].double().numpy(),
                             dx=it["dx"].double().numpy())
        def row(q, floor=0.0):
            def c(a, b):
                if q not in outs[a] or q not in outs[b]:
                    return "-"
                return f"{rel(outs[a][q], outs[b][q], floor):.1e}"
            print(f"| {name} | {q} | {c('ref1', 'o64')} | {c('refN', 'ref1')} | {c('o32', 'o64')} | {c('o32', 'ref1')} | {c('emu3', 'o64')} |", flush=True)
        row("t")
        row("code", 0.05)
        row("b")
        row("dx")


if __name__ == "__main__":
    main()
