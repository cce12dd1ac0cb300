"""TEST INFRASTRUCTURE (CPU only): VERDICT r1 item 4(iii) - a two-product reverse sweep for the SURFACE rows only.

Emulates k_decoder_p3 (scripts/tmem_rz_emulation.py: fp16 pairs, three products, accumulator truncated per MMA step) and
two cheaper variants in which the reverse sweep of the surface-row call (the M rows of the SDF term) drops one product:
  adj   G_hi W_hi + G_hi W_lo     adjoints rounded to fp16, weights exact    (per-row random error)
  wgt   G_hi W_hi + G_lo W_hi     weights rounded to fp16, adjoints exact    (the same perturbed network for every row)
For N objects of kitti_batch32 from the fp32 oracle's states before iterations 0, 4, 9: one GN iteration on the emulated
decoder against the fp64 oracle - new code / pose / dx, and the Jacobian rows of the surface term (maximum over rows that
are not ReLU flips in the three-product run).  usage: python scripts/two_product_study.py [N]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import torch

from qsp_slam_b200 import synth
from oracle import gn_oracle as G
import noise_floor as NF
import tmem_rz_emulation as RZ

VARIANT = {"v": "p3", "M": -1}


def acc_mm(a, w, sa, sw, drop):
    a = a.float().double()
    w = w.float().double()
    ah, al, wh, wl = NF.hi(a * sa), NF.lo(a * sa), NF.hi(w * sw), NF.lo(w * sw)
    K = a.shape[1]
    large = torch.zeros(a.shape[0], w.shape[0], dtype=torch.float64)
    small = torch.zeros_like(large)
    for kb in range(0, K, 64):
        for k0 in range(kb, min(kb + 64, K), 16):
            sl = slice(k0, min(k0 + 16, K))
            large = RZ.rz32(large + ah[:, sl] @ wh[:, sl].t())
            if drop != "adj":
                small = RZ.rz32(small + al[:, sl] @ wh[:, sl].t())
        if drop != "wgt":
            for k0 in range(kb, min(kb + 64, K), 16):
                sl = slice(k0, min(k0 + 16, K))
                small = RZ.rz32(small + ah[:, sl] @ wl[:, sl].t())
    return RZ.rn32(large + small) / (sa * sw)


def emu(dec, code, x):
    n = x.shape[0]
    drop = VARIANT["v"] if n == VARIANT["M"] else "p3"
    inp = torch.cat([code.double().expand(n, -1), x.double()], 1)
    a = inp
    masks = []
    for l in range(8):
        if l == 4:
            a = torch.cat([a, inp], 1)
        a = RZ.rn32(acc_mm(a, dec.W[l], 16.0, 16.0, "p3") + dec.b[l])
        masks.append(a > 0)
        a = torch.relu(a)
    s = torch.tanh(RZ.rn32(a @ dec.W[8][0] + dec.b[8][0]))
    g = ((1 - s * s)[:, None] * dec.W[8]) * masks[7]
    skip = None
    for l in range(7, 0, -1):
        g = acc_mm(g, dec.W[l].t(), 256.0, 16.0, drop)
        if l == 4:
            skip = g[:, -67:]
            g = g[:, :-67]
        g = RZ.rn32(g * masks[l - 1])
    g = acc_mm(g, dec.W[0].t(), 256.0, 16.0, drop) + skip
    return s, RZ.rn32(g)


def iteration(fd, prm, t_oc, z, o, variant):
    dt = torch.float64
    args = (fd, prm, t_oc.to(dt), z.to(dt), torch.from_numpy(o["pts"]).to(dt), torch.from_numpy(o["rays"]).to(dt), torch.from_numpy(o["depth"]).to(dt))
    if variant is None:
        return G.gn_iteration(*args)
    VARIANT["v"], VARIANT["M"] = variant, o["pts"].shape[0]
    keep = G.decoder_forward_jacobian
    G.decoder_forward_jacobian = emu
    try:
        return G.gn_iteration(*args)
    finally:
        G.decoder_forward_jacobian = keep


def main():
    n_obj = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    torch.set_num_threads(os.cpu_count() or 1)
    state = synth.make_decoder_state(0)
    fd32, fd64 = G.FoldedDecoder(state), G.FoldedDecoder(state, torch.float64)
    cfg, objs = synth.make_workload("kitti_batch32", state, objects=n_obj)
    prm = G.Params(cfg)
    print("| object | it | variant | code | pose | dx | b | surface rows: median | p99 | max (non-flip) | rows > 1e-3 |")
    print("|---|---|---|---|---|---|---|---|---|---|---|")
    worst = {}
    for oi, o in enumerate(objs):
        tr = []
        G.reconstruct_object(fd32, prm, o["t_cam_obj"], o["pts"], o["rays"], o["depth"], trace=tr)
        for e in (0, 4, 9):
            if len(tr) <= e:
                continue
            t_oc = torch.inverse(torch.from_numpy(o["t_cam_obj"].copy())) if e == 0 else tr[e - 1]["t_obj_cam"]
            z = torch.zeros(64) if e == 0 else tr[e - 1]["code"]
            i64 = iteration(fd64, prm, t_oc.double(), z, o, None)
            g64 = i64["sdf"]["grad"]
            flip3 = None
            for v in ("p3", "adj", "wgt"):
                ie = iteration(fd64, prm, t_oc.double(), z, o, v)
                r = lambda q, fl=0.0: NF.rel(ie[q].numpy(), i64[q].numpy(), fl)
                rows = ((ie["sdf"]["grad"] - g64).norm(dim=1) / g64.norm(dim=1)).numpy()
                if v == "p3":
                    flip3 = rows > 1e-3
                nf = rows[~flip3]
                w = worst.setdefault(v, dict(code=0, pose=0, dx=0, rowmax=0, over=0, rows=0))
                w["code"] = max(w["code"], r("code", 0.05)); w["pose"] = max(w["pose"], r("t_obj_cam")); w["dx"] = max(w["dx"], r("dx"))
                w["rowmax"] = max(w["rowmax"], float(nf.max())); w["over"] += int((nf > 1e-3).sum()); w["rows"] += nf.size
                print(f"| {oi} | {e} | {v} | {r('code', 0.05):.1e} | {r('t_obj_cam'):.1e} | {r('dx'):.1e} | {r('b'):.1e} | "
                      f"{np.median(rows):.1e} | {np.quantile(rows, 0.99):.1e} | {nf.max():.1e} | {int((nf > 1e-3).sum())} |", flush=True)
    print("\n| variant | worst code | worst pose | worst dx | max row error (non-flip) | rows > 1e-3 | rows |")
    print("|---|---|---|---|---|---|---|")
    for v, w in worst.items():
        print(f"| {v} | {w['code']:.1e} | {w['pose']:.1e} | {w['dx']:.1e} | {w['rowmax']:.1e} | {w['over']} | {w['rows']} |")


if __name__ == "__main__":
    main()
