"""CPU study (oracle only): how many first-pass rows a front-to-back phase schedule decodes per GN iteration.

The first pass decodes a ray's in-ball samples in phases and stops at the first
certainly-opaque one.  Any schedule gives the same result; what changes is the number of rows decoded BEHIND that
sample (wasted) and the number of launches.  Compared here on the oracle's own trajectories:
  fixed32    phases of 32 samples (the r01 scheme)
  hint+s     phase 0 = [0, h + s) with h = the ray's first-opaque position of the PREVIOUS iteration (32 in iteration 0),
             later phases of `mid` samples, the last phase takes the rest
  ideal      everything up to and including the first opaque sample
usage: python scripts/phase_schedule_sim.py [kitti_batch32|kitti_single|redwood_chair] [objects]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from qsp_slam_b200 import synth
from oracle import gn_oracle as G


def per_ray(ren, R, lim):
    """in-ball count and position (count of samples up to and including it) of the first certainly-opaque sample."""
    vr = ren["valid_ray"].numpy()
    s = ren["sdf"].numpy()
    n = np.bincount(vr, minlength=R)
    first = n.copy()                                       # no opaque sample: everything is needed
    start = np.concatenate([[0], np.cumsum(n)])
    opq = s <= -lim
    idx = np.nonzero(opq)[0]
    if idx.size:
        r = vr[idx]
        pos = idx - start[r] + 1
        # first occurrence per ray
        order = np.lexsort((pos, r))
        r_s, pos_s = r[order], pos[order]
        keep = np.concatenate([[True], r_s[1:] != r_s[:-1]])
        first[r_s[keep]] = pos_s[keep]
    return n, first


def rows_sched(n, first, lens_of):
    """lens_of(ray index, phase) -> length; the last phase takes the rest.  Returns total rows and rows per phase."""
    tot = 0
    for i, (nn, f) in enumerate(zip(n, first)):
        done = 0
        p = 0
        while done < nn:
            L = lens_of(i, p)
            m = (nn - done) if L is None else min(L, nn - done)
            done += m
            p += 1
            if f < nn and done >= f:
                break
        tot += done
    return tot


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "kitti_batch32"
    n_obj = int(sys.argv[2]) if len(sys.argv) > 2 else (4 if wl == "kitti_batch32" else 1)
    state = synth.make_decoder_state(0)
    fd = G.FoldedDecoder(state)
    cfg, objs = synth.make_workload(wl, state, objects=n_obj)
    prm = G.Params(cfg)
    S = prm.S
    n_ph = (S + 31) // 32
    if n_ph > 1 and S - (n_ph - 1) * 32 <= 8:
        n_ph -= 1
    torch.set_num_threads(os.cpu_count() or 1)
    print(f"# {wl}: S = {S}, fixed scheme runs {n_ph} phases")
    print("| object | it | in-ball | ideal | fixed32 | hint+1 | hint+2 | hint+3 | hint+4 | hint+2, mid 8 |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    tot = np.zeros(8)
    for oi, o in enumerate(objs):
        tr = []
        G.reconstruct_object(fd, prm, o["t_cam_obj"], o["pts"], o["rays"], o["depth"], trace=tr)
        R = o["rays"].shape[0]
        hint = np.full(R, 32)
        for e, it in enumerate(tr):
            code_norm = float(torch.linalg.norm(it["code_in"])) if "code_in" in it else 0.5
            lim = prm.th + 0.0045 + 0.004 * code_norm
            n, first = per_ray(it["render"], R, lim)
            ideal = int(np.minimum(first, n).sum())
            fixed = rows_sched(n, first, lambda i, p: None if p == n_ph - 1 else 32)
            cols = []
            for s_, mid in ((1, 32), (2, 32), (3, 32), (4, 32), (2, 8)):
                cols.append(rows_sched(n, first, lambda i, p: None if p == n_ph - 1 else (hint[i] + s_ if p == 0 else mid)))
            row = [int(n.sum()), ideal, fixed] + cols
            tot += np.array(row, float)
            print(f"| {oi} | {e} | " + " | ".join(str(v) for v in row) + " |", flush=True)
            hint = np.where(first < n, first, n)
    print("| all | | " + " | ".join(f"{int(v)}" for v in tot) + " |")
    print("| relative to fixed32 | | " + " | ".join(f"{v / tot[2]:.3f}" for v in tot) + " |")


if __name__ == "__main__":
    main()
